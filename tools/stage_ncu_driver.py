#!/usr/bin/env python
"""Launches each stage kernel a few times on 64 1080p frames (for an ncu capture of the stage kernels:
ncu --set full -k regex:'stencil3|ssim_moments|overlay|front|expand' ... python tools/stage_ncu_driver.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import norm_b200 as nb

W, H, n = 1920, 1080, 64
src_t = torch.randint(0, 256, (n, H, W, 4), dtype=torch.uint8, device="cuda")
dst_t = torch.empty_like(src_t)
torch.cuda.synchronize()
ctx = nb.Context(0)
src = ctx.image_wrap(src_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
dst = ctx.image_wrap(dst_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
rng = np.random.default_rng(2)
circles = [nb.Circle((float(rng.uniform(0, 1)), float(rng.uniform(0, 1))), float(rng.uniform(0.005, 0.05))) for _ in range(64)]
draw = nb.DrawUniform()
draw.toggle_crosshair(), draw.set_circle_center(960.0, 540.0), draw.set_circle_radius(25.0)
fe = nb.FeatureUniform(W, H)
fe.raw.edge_detect = 1
for _ in range(2):
    ctx.sobel(src, dst)
    ctx.sharpen(16, src, dst)
    ctx.box_blur(src, dst)
    ctx.ssim(src, dst)
    ctx.overlay(src, dst, circles, 5, draw)
    ctx.effects(fe, src, dst, nb.NORM_CS_REFERENCE)
ctx.sync()
print("ok")
