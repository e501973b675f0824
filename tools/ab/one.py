#!/usr/bin/env python
"""Times a fixed list of chain / blur configurations with ONE build of the library (A/B helper, one process).
    python tools/ab/one.py --lib tools/ab/x.so [--frames 256] [--reps 5] [--set chain|all]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import norm_b200 as nb

ap = argparse.ArgumentParser()
ap.add_argument("--lib", default=None)
ap.add_argument("--frames", type=int, default=256)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--set", default="chain")
ap.add_argument("--tag", default=None)
a = ap.parse_args()
if a.lib:
    nb._lib.LIB_PATH = os.path.abspath(a.lib)
tag = a.tag or (os.path.basename(a.lib).replace(".so", "") if a.lib else "tree")
W, H, n = 1920, 1080, a.frames
src_t = torch.randint(0, 256, (n, H, W, 4), dtype=torch.uint8, device="cuda")
dst_t = torch.empty_like(src_t)
ss = torch.zeros(n, dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
ctx = nb.Context(0)
src = ctx.image_wrap(src_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
dst = ctx.image_wrap(dst_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
cfgs = [("chain", 21, 0), ("chain", 7, 0), ("chain", 3, 0), ("chain", 21, 1), ("chain", 3, 1)]
if a.set == "all":
    cfgs += [("chain", 39, 0), ("chain", 39, 1), ("blur", 21, 0), ("blur", 7, 0), ("blur", 21, 1), ("blur", 39, 0)]
out = []
for what, r, gray in cfgs:
    f = nb.FeatureUniform(W, H)
    f.raw.blur, f.raw.blur_radius, f.raw.edge_detect, f.raw.grayscale = 1, r, 1 if what == "chain" else 0, gray
    if what == "chain":
        run = lambda: ctx.chain_async(f, src, dst, nb.NORM_CS_RAW, ss.data_ptr())
    else:
        run = lambda: ctx.effects(f, src, dst, nb.NORM_CS_RAW)
    best = 1e9
    for rep in range(2):
        run()
        ctx.sync()
        ctx.timer_start()
        for _ in range(a.reps):
            run()
        ctx.timer_stop()
        best = min(best, ctx.timer_elapsed_ms() / a.reps)
    out.append("%s%s r=%d: %.1f" % (what, "-gray" if gray else "", r, n * W * H / best / 1e6))
print("%-10s" % tag, " | ".join(out), "Gpx/s")
