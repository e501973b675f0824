#!/usr/bin/env python
"""Times every single-stage entry point on a 1080p batch and reports algorithmic GB/s against the
measured HBM peak (MEASURED_PEAKS.json).  Stage kernels are HBM-bound: this is their roofline check."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import norm_b200 as nb

if os.environ.get("NORM_LIB"):  # A/B against another build of the library
    nb._lib.LIB_PATH = os.path.abspath(os.environ["NORM_LIB"])
W, H, n = 1920, 1080, int(os.environ.get("FRAMES", 256))
peak = 6551
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
src_t = torch.randint(0, 256, (n, H, W, 4), dtype=torch.uint8, device="cuda")
dst_t = torch.empty_like(src_t)
torch.cuda.synchronize()
ctx = nb.Context(0)
src = ctx.image_wrap(src_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
dst = ctx.image_wrap(dst_t.data_ptr(), W, H, n, W * 4, W * 4 * H)
px = n * W * H


def timeit(name, fn, bytes_per_px, reps=5):
    fn()
    ctx.timer_start()
    for _ in range(reps):
        fn()
    ctx.timer_stop()
    ms = ctx.timer_elapsed_ms() / reps
    gbs = px * bytes_per_px / ms / 1e6
    print(f"{name:34s} {ms:8.3f} ms  {px/ms/1e6:8.1f} Gpx/s  {gbs:8.0f} GB/s  {gbs/peak*100:5.1f}% of measured HBM peak")
    return gbs


f = nb.FeatureUniform(W, H)
f.raw.grayscale = 1
timeit("front: grayscale (raw)", lambda: ctx.front(f, src, dst, nb.NORM_CS_RAW), 8)
timeit("front: sRGB table + grayscale", lambda: ctx.front(f, src, dst, nb.NORM_CS_REFERENCE), 8)
f2 = nb.FeatureUniform(W, H)
f2.raw.invert = 1
timeit("front: invert", lambda: ctx.front(f2, src, dst, nb.NORM_CS_RAW), 8)
timeit("sobel", lambda: ctx.sobel(src, dst), 8)
timeit("sharpen", lambda: ctx.sharpen(16, src, dst), 8)
timeit("ssim (two images)", lambda: ctx.ssim(src, dst), 8)
fe = nb.FeatureUniform(W, H)
fe.raw.edge_detect = 1
timeit("effects: sRGB table + edge (1 frame pass)", lambda: ctx.effects(fe, src, dst, nb.NORM_CS_REFERENCE), 8)
fe.raw.grayscale = 1
timeit("effects: sRGB + gray + edge", lambda: ctx.effects(fe, src, dst, nb.NORM_CS_REFERENCE), 8)
fs = nb.FeatureUniform(W, H)
fs.raw.sharpen = 1
timeit("effects: sRGB table + sharpen", lambda: ctx.effects(fs, src, dst, nb.NORM_CS_REFERENCE), 8)
timeit("box blur 3x3", lambda: ctx.box_blur(src, dst), 8)
rng = np.random.default_rng(2)
circles = [nb.Circle((float(rng.uniform(0, 1)), float(rng.uniform(0, 1))), float(rng.uniform(0.005, 0.05))) for _ in range(64)]
draw = nb.DrawUniform()
draw.toggle_crosshair(), draw.set_circle_center(960.0, 540.0), draw.set_circle_radius(25.0)
timeit("overlay: no shapes (copy)", lambda: ctx.overlay(src, dst, []), 8)
timeit("overlay: 64 circles + crosshair", lambda: ctx.overlay(src, dst, circles, 5, draw), 8)
big = [nb.Circle((0.5, 0.5), 0.45)] + circles
timeit("overlay: + one circle over 1/3 frame", lambda: ctx.overlay(src, dst, big, 0, draw), 8)
for r in (3, 7, 21):
    timeit(f"blur r={r} (RGBA, 4 channels)", lambda: ctx.blur(r, src, dst), 8)
# expansion: one 7680x4320 image per call, RGB8 -> RGBA8 (7 B/px) and 1-bit gray (4.125 B/px); includes the H2D of the packed scanlines
w8, h8 = 7680, 4320
out8 = ctx.image_create(w8, h8, 1)
rgb = np.random.default_rng(0).integers(0, 256, size=w8 * h8 * 3, dtype=np.uint8)
png = nb.Png(w8, h8, 8, 2, rgb)
px = w8 * h8
timeit("expand RGB8 8K (incl. H2D pageable)", lambda: ctx.expand_rgba8(png, nb.NORM_EXPAND_SPEC, out8, 0), 7, reps=3)
# the expansion kernel itself: packed scanlines already in HBM (16-byte row pitch), 4 frames of 7680x4320 per launch,
# against in_bpp/8 + 4 algorithmic bytes per pixel
nf = 4
out8b = ctx.image_create(w8, h8, nf)
px = w8 * h8 * nf
CH = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}
for (name, ct, bd, mode) in (("RGBA8", 6, 8, nb.NORM_EXPAND_SPEC), ("RGB8", 2, 8, nb.NORM_EXPAND_SPEC), ("RGB8 compat", 2, 8, nb.NORM_EXPAND_COMPAT),
                             ("gray8", 0, 8, nb.NORM_EXPAND_SPEC), ("gray+alpha8", 4, 8, nb.NORM_EXPAND_SPEC),
                             ("palette8 + tRNS", 3, 8, nb.NORM_EXPAND_SPEC), ("gray4", 0, 4, nb.NORM_EXPAND_SPEC),
                             ("palette2", 3, 2, nb.NORM_EXPAND_SPEC), ("gray1", 0, 1, nb.NORM_EXPAND_SPEC),
                             ("gray16", 0, 16, nb.NORM_EXPAND_SPEC), ("RGB16", 2, 16, nb.NORM_EXPAND_SPEC), ("RGBA16", 6, 16, nb.NORM_EXPAND_SPEC)):
    rowb8 = (w8 * CH[ct] * bd + 7) // 8
    pitch8 = (rowb8 + 15) // 16 * 16
    fstride = pitch8 * h8 + 256
    buf = torch.randint(0, 256, (nf * fstride + 256,), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    meta = nb.Png(w8, h8, bd, ct, np.zeros(1, np.uint8), plte=np.arange(768, dtype=np.uint8) if ct == 3 else None,
                  trns=np.arange(200, dtype=np.uint8) if ct == 3 else None)
    timeit(f"expand {name} 8K x{nf}, device-resident", lambda: ctx.expand_rgba8_device(meta, buf.data_ptr(), pitch8, fstride, nf, mode, out8b, 0),
           CH[ct] * bd / 8 + 4, reps=5)
    del buf
out8b.destroy()
px = w8 * h8

# scanline unfilter (GPU wavefront) vs the host: a 1920x1080 RGB8 stream with all five filter types
import time
import zlib
rng = np.random.default_rng(1)
wu, hu = 1920, 1080
rowb = wu * 3
stream = bytearray()
for y in range(hu):
    stream.append(y % 5)
    stream += rng.integers(0, 256, size=rowb, dtype=np.uint8).tobytes()
stream = bytes(stream)
ctx.unfilter(wu, hu, 8, 2, 0, stream)
t0 = time.perf_counter()
for _ in range(5):
    ctx.unfilter(wu, hu, 8, 2, 0, stream)
dt = (time.perf_counter() - t0) / 5
print(f"{'unfilter 1080p RGB8 (1 CTA wavefront, incl. H2D+D2H)':54s} {dt*1e3:8.3f} ms  {len(stream)/dt/1e9:6.2f} GB/s of filtered bytes")
