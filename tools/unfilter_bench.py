"""Throughput of the GPU scanline unfilter + expansion: one image vs a batch.
Streams come from the GPU filter selection (mixed filter types, as an encoder would write them).
Times the whole call: H2D of the inflated streams, unfilter, expansion into device frames."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import norm_b200 as nb  # noqa: E402
from conftest import smooth_rgba  # noqa: E402


def main():
    w, h, distinct = 1920, 1080, 8
    ctx = nb.Context(0)
    rng = np.random.default_rng(1)
    streams = []
    for s in range(distinct):
        f = (smooth_rgba(w, h, seed=s)[..., :3] + rng.integers(0, 3, size=(h, w, 3))).astype(np.uint8)
        st, ft = ctx.filter_scanlines(w, h, 8, 2, f)
        streams.append(st.tobytes())
    hist = np.bincount(ft, minlength=5).tolist()
    rows = []
    import ctypes as C
    from norm_b200._lib import NormInflatedDesc
    slen = len(streams[0])
    pinned, handle = ctx.host_alloc(slen * distinct)  # the streams in pinned memory: H2D at PCIe speed
    for i, st in enumerate(streams):
        pinned[i * slen:(i + 1) * slen] = np.frombuffer(st, np.uint8)
    for n in (1, 16, 148, 296, 592, 1184):
        out = ctx.image_create(w, h, n)
        descs = (NormInflatedDesc * n)()
        for i in range(n):
            d = descs[i]
            d.width, d.height, d.bit_depth, d.color_type, d.interlace = w, h, 8, 2, 0
            d.data, d.len = pinned.ctypes.data + (i % distinct) * slen, slen

        def call():
            ctx._ok(ctx.lib.norm_cuda_unfilter_expand_rgba8_batch(ctx.h, descs, n, None, 0, None, 0, nb.NORM_EXPAND_COMPAT,
                                                                 out.handle, 0))
        call()  # warm-up, sizes scratch
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter()
            call()
            best = min(best, time.perf_counter() - t0)
        rows.append({"images": n, "ms": best * 1e3, "ms_per_image": best * 1e3 / n,
                     "Mpixel/s": n * w * h / best / 1e6, "stream_GB/s": n * len(streams[0]) / best / 1e9})
        out.destroy()
        print(rows[-1], file=sys.stderr)
    print(json.dumps({"workload": "unfilter+expand, 1920x1080 RGB8, filters chosen by the encoder heuristic",
                      "filter_histogram_last_frame": hist, "timing": "host wall clock around the C-ABI call; streams in pinned host memory, H2D inside the call",
                      "rows": rows}))


if __name__ == "__main__":
    main()
