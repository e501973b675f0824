#!/usr/bin/env python
"""Small driver for ncu / quick timing: runs the chain on N 1080p frames (or one WxH image)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import norm_b200 as nb

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=64)
ap.add_argument("--radius", type=int, default=21)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--gray", type=int, default=0)
ap.add_argument("--no-ssim", action="store_true")
ap.add_argument("--blur-only", action="store_true", help="time norm_cuda_blur (all four channels) instead of the chain")
ap.add_argument("--lib", default=None, help="alternative libnorm_cuda.so (A/B runs)")
a = ap.parse_args()
if a.lib:
    nb._lib.LIB_PATH = os.path.abspath(a.lib)
W, H, n = a.width, a.height, a.frames
src_t = torch.randint(0, 256, (n, H, W, 4), dtype=torch.uint8, device="cuda")
dst_t = torch.empty_like(src_t)
ss = torch.zeros(n, dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
ctx = nb.Context(0)
src = ctx.image_wrap(src_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
dst = ctx.image_wrap(dst_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
f = nb.FeatureUniform(W, H)
f.raw.blur, f.raw.blur_radius, f.raw.edge_detect, f.raw.grayscale = 1, a.radius, 1, a.gray
if a.blur_only:
    f.raw.edge_detect = 0
    run = lambda: ctx.effects(f, src, dst, nb.NORM_CS_RAW)
else:
    run = lambda: ctx.chain_async(f, src, dst, nb.NORM_CS_RAW, None if a.no_ssim else ss.data_ptr())
run()
ctx.sync()
ctx.timer_start()
for _ in range(a.reps):
    run()
ctx.timer_stop()
ms = ctx.timer_elapsed_ms() / a.reps
what = "blur-only" if a.blur_only else "chain"
print(f"{what} r={a.radius} gray={a.gray} {n}x{W}x{H}: {ms:.3f} ms/pass  {n*W*H/ms/1e6:.1f} Gpx/s  {n*W*H*8/ms/1e6:.0f} GB/s algorithmic")
