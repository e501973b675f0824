#!/usr/bin/env python
"""Chain throughput (blur r=21 + Sobel + SSIM) of ONE image for several geometries: python tools/ab/geom.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import norm_b200 as nb
ctx = nb.Context(0)
GEOMS = [(32768, 4118), (32640, 4118), (35520, 4118), (32768, 2048), (16384, 8192), (7680, 17000), (1920, 69120), (32768, 16384)]
if len(sys.argv) > 1:
    GEOMS = [tuple(int(v) for v in a.split('x')) for a in sys.argv[1:]]
for (W, H) in GEOMS:
    src_t = torch.randint(0, 256, (1, H, W, 4), dtype=torch.uint8, device="cuda")
    dst_t = torch.empty_like(src_t)
    ss = torch.zeros(1, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    src = ctx.image_wrap(src_t.data_ptr(), W, H, 1, W * 4, W * 4 * H)
    dst = ctx.image_wrap(dst_t.data_ptr(), W, H, 1, W * 4, W * 4 * H)
    f = nb.FeatureUniform(W, H); f.raw.blur, f.raw.blur_radius, f.raw.edge_detect = 1, 21, 1
    run = lambda: ctx.chain_async(f, src, dst, nb.NORM_CS_RAW, ss.data_ptr())
    run(); ctx.sync(); ctx.timer_start()
    for _ in range(5): run()
    ctx.timer_stop(); ms = ctx.timer_elapsed_ms() / 5
    strips = (W + 119) // 120
    print(f"bands={os.environ.get('NORM_CUDA_BANDS', 'auto')} {W}x{H}: {ms:.3f} ms  {W*H/ms/1e6:.1f} Gpx/s  strips {strips}")
    src.destroy(); dst.destroy(); del src_t, dst_t
