"""GPU scanline filter selection for the encoder against the CPU oracle, plus round trips
through the GPU unfilter and through PIL."""
import io
import os
import struct
import zlib

import numpy as np
import pytest

import norm_b200 as nb
from conftest import GOLDEN, smooth_rgba, splitmix_rgba

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = nb.Context(0)
    yield c
    c.close()


def png_from_stream(w, h, bd, ct, stream):
    def chunk(typ, body):
        return struct.pack(">I", len(body)) + typ + body + struct.pack(">I", zlib.crc32(typ + body) & 0xFFFFFFFF)
    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, bd, ct, 0, 0, 0)) +
            chunk(b"IDAT", zlib.compress(bytes(stream), 1)) + chunk(b"IEND", b""))


@pytest.mark.parametrize("compat", [0, 1])
def test_filter_matches_oracle(ctx, oracle, compat):
    rng = np.random.default_rng(11)
    for (w, h, ch, bd, ct) in ((61, 40, 4, 8, 6), (333, 70, 3, 8, 2), (50, 9, 1, 8, 0), (20, 11, 2, 16, 4), (7, 5, 4, 16, 6),
                               (1, 1, 4, 8, 6), (130, 3, 1, 2, 0)):
        bpp, rowb = max(1, ch * bd // 8), (w * ch * bd + 7) // 8
        base = smooth_rgba(rowb, h, seed=w)[..., 0]
        px = (base + rng.integers(0, 3, size=base.shape)).astype(np.uint8)
        es, ef = oracle.filter_scanlines(px, h, rowb, bpp, compat)
        gs, gf = ctx.filter_scanlines(w, h, bd, ct, px, compat)
        assert np.array_equal(gf, ef) and np.array_equal(gs, es), (w, h, ch, bd)
        if compat == 0:  # what the GPU filtered, the GPU unfilters
            assert np.array_equal(ctx.unfilter(w, h, bd, ct, 0, gs.tobytes()), px.reshape(-1))


def test_filter_obama_and_all_types_used(ctx, oracle):
    d = oracle.png_decode(open(os.path.join(GOLDEN, "images", "obama.png"), "rb").read())
    for compat in (0, 1):
        es, ef = oracle.filter_scanlines(d.pixels, d.height, d.width * 3, 3, compat)
        gs, gf = ctx.filter_scanlines(d.width, d.height, 8, 2, d.pixels, compat)
        assert np.array_equal(gs, es) and np.array_equal(gf, ef)
    assert len(set(gf.tolist())) >= 3  # a photograph exercises several filter types


def test_effect_result_leaves_as_png(ctx, oracle):
    """chain on the GPU -> filter_image on the device-resident result -> deflate on the host -> PIL decodes the
    same RGBA8 the GPU produced."""
    from PIL import Image
    w, h = 200, 120
    img = smooth_rgba(w, h, seed=2)
    f = nb.FeatureUniform(w, h)
    f.toggle_blur(), f.toggle_edge_detect()
    f.raw.blur_radius = 7
    src, dst = ctx.image_from_host(img), ctx.image_create(w, h, 1)
    ctx.chain(f, src, dst, nb.NORM_CS_RAW)
    result = dst.download()[0]
    stream, ft = ctx.filter_image(dst)
    es, ef = oracle.filter_scanlines(result, h, w * 4, 4, 0)
    assert np.array_equal(stream, es) and np.array_equal(ft, ef)
    back = np.asarray(Image.open(io.BytesIO(png_from_stream(w, h, 8, 6, stream))))
    assert np.array_equal(back, result)
    src.destroy(), dst.destroy()


def test_filter_large_frame(ctx, oracle):
    w, h = 1920, 1080
    img = splitmix_rgba(w, h, seed=3) // 4 + smooth_rgba(w, h, seed=4) // 2
    d = ctx.image_from_host(img)
    stream, ft = ctx.filter_image(d)
    es, ef = oracle.filter_scanlines(img, h, w * 4, 4, 0)
    assert np.array_equal(stream, es) and np.array_equal(ft, ef)
    d.destroy()
