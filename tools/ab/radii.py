#!/usr/bin/env python
"""Chain / blur throughput for a list of radii (A/B helper): python tools/ab/radii.py 23 25 27 29"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import norm_b200 as nb
W, H, n = 1920, 1080, 128
src_t = torch.randint(0, 256, (n, H, W, 4), dtype=torch.uint8, device="cuda")
dst_t = torch.empty_like(src_t)
ss = torch.zeros(n, dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
ctx = nb.Context(0)
src = ctx.image_wrap(src_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
dst = ctx.image_wrap(dst_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
out = []
for r in [int(x) for x in sys.argv[1:]]:
    for what in ("chain", "blur"):
        f = nb.FeatureUniform(W, H)
        f.raw.blur, f.raw.blur_radius, f.raw.edge_detect = 1, r, 1 if what == "chain" else 0
        run = (lambda: ctx.chain_async(f, src, dst, nb.NORM_CS_RAW, ss.data_ptr())) if what == "chain" else (lambda: ctx.effects(f, src, dst, nb.NORM_CS_RAW))
        run(); ctx.sync(); ctx.timer_start()
        for _ in range(3): run()
        ctx.timer_stop()
        out.append("%s r=%d: %.1f" % (what, r, n * W * H / (ctx.timer_elapsed_ms() / 3) / 1e6))
print(os.environ.get("NORM_CUDA_FUSED_GEN", "gen2"), " | ".join(out))
