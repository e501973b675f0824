#!/usr/bin/env python
"""Cost of the band scheduler's three-launch step against one launch over the same extended band, on one GPU:
three emulated ranks of a 32768 x 12288 image (4096 rows each); the middle rank has both halos."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import norm_b200 as nb
W, H, N = 32768, 12288, 3
f = nb.FeatureUniform(W, H); f.raw.blur, f.raw.blur_radius, f.raw.edge_detect = 1, 21, 1
ctxs = [nb.Context(0) for _ in range(N)]
bands = [nb.Band(c, f, W, H, N, r) for r, c in enumerate(ctxs)]
hs = [b.export() for b in bands]
for b in bands:
    b.connect(hs)
    t = torch.randint(0, 256, (b.src.height, W, 4), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    b.ctx.copy_rows(b.src, 0, t.data_ptr(), W * 4, 0, b.src.height); b.ctx.sync(); del t
    b.publish(); b.ctx.sync()
mid, c = bands[1], ctxs[1]
mom = c.image_create(64, 1, 1)
def timeit(fn, reps=10):
    fn(); mid.sync(); c.timer_start()
    for _ in range(reps): fn()
    c.timer_stop(); mid.sync()
    return c.timer_elapsed_ms() / reps
t_split = timeit(lambda: mid.step_async(nb.NORM_CS_RAW, mom.device_ptr))
fb = nb.FeatureUniform(W, mid.src.height); fb.raw.blur, fb.raw.blur_radius, fb.raw.edge_detect = 1, 21, 1
t_one = timeit(lambda: c.chain_band_async(fb, mid.src, mid.dst, mid.halo_up, mid.rows, nb.NORM_CS_RAW, mom.device_ptr))
px = W * mid.rows
print(f"band step (3 launches + halo copies): {t_split:.3f} ms = {px/t_split/1e6:.1f} Gpx/s;  one launch over band+halo: {t_one:.3f} ms = {px/t_one/1e6:.1f} Gpx/s")
