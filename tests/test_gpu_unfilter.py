"""GPU scanline unfilter + Adam7 de-interlace against the reference's golden blobs
and the CPU oracle: bit-exact pixel_buffer for every valid PngSuite file and tests/obama.png."""
import glob
import os
import struct
import zlib

import numpy as np
import pytest

import norm_b200 as nb
from conftest import GOLDEN, PNGSUITE

pytestmark = pytest.mark.gpu

VALID = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(PNGSUITE, "*.png"))
               if not os.path.basename(p).startswith("x"))
BLOBS = sorted(os.path.basename(p)[:-5] for p in glob.glob(os.path.join(PNGSUITE, "*.blob")))


def inflate_png(path):
    """Chunk walk + zlib inflate on the host (stays on the CPU: serial Huffman decoding).
    -> (width, height, bit_depth, color_type, interlace, inflated bytes, plte, trns)"""
    data = open(path, "rb").read()
    pos, idat, plte, trns, hdr = 8, b"", b"", b"", None
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if typ == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        elif typ == b"IDAT":
            idat += body
        elif typ == b"PLTE":
            plte = body
        elif typ == b"tRNS":
            trns = body
        elif typ == b"IEND":
            break
        pos += 12 + n
    w, h, bd, ct, _, _, il = hdr
    return w, h, bd, ct, il, zlib.decompress(idat), np.frombuffer(plte, np.uint8), np.frombuffer(trns, np.uint8)


@pytest.fixture(scope="module")
def ctx():
    c = nb.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", BLOBS)
def test_unfilter_matches_reference_blob(ctx, name):
    """The 13 blobs the reference's own tests pin (src/png/decoder.rs:295-365)."""
    w, h, bd, ct, il, raw, _, _ = inflate_png(os.path.join(PNGSUITE, name + ".png"))
    blob = open(os.path.join(PNGSUITE, name + ".blob"), "rb").read()
    assert ctx.unfilter(w, h, bd, ct, il, raw).tobytes() == blob[17:]


@pytest.mark.parametrize("name", VALID)
def test_unfilter_pngsuite(ctx, oracle, name):
    path = os.path.join(PNGSUITE, name + ".png")
    w, h, bd, ct, il, raw, plte, trns = inflate_png(path)
    d = oracle.png_decode(open(path, "rb").read())
    got = ctx.unfilter(w, h, bd, ct, il, raw)
    assert np.array_equal(got, d.pixels), name
    # inflated stream -> RGBA8 without leaving the device
    out = ctx.image_create(w, h, 1)
    ctx.unfilter_expand(w, h, bd, ct, il, raw, nb.NORM_EXPAND_SPEC, out, 0, plte, trns)
    assert np.array_equal(out.download()[0], oracle.expand_rgba8(d, oracle.EXPAND_SPEC)), name
    out.destroy()


def test_unfilter_obama_and_large_rows(ctx, oracle):
    """683 rows > one block of 512 threads: exercises the hand-over between row blocks."""
    path = os.path.join(GOLDEN, "images", "obama.png")
    w, h, bd, ct, il, raw, _, _ = inflate_png(path)
    d = oracle.png_decode(open(path, "rb").read())
    assert np.array_equal(ctx.unfilter(w, h, bd, ct, il, raw), d.pixels)


def test_unfilter_synthetic_all_filters(ctx):
    """Random data pushed through a host-side PNG filter with every filter type (incl. rows > 512 and
    16-bit RGBA, bpp = 8), then unfiltered on the GPU: round trip must be the identity."""
    rng = np.random.default_rng(3)
    for (w, h, ch, bd) in ((37, 600, 3, 8), (19, 33, 4, 16), (300, 20, 1, 8), (65, 1100, 2, 8)):
        bpp = max(1, ch * bd // 8)
        rowb = (w * ch * bd + 7) // 8
        px = rng.integers(0, 256, size=(h, rowb), dtype=np.uint8)
        stream = bytearray()
        prev = np.zeros(rowb, np.int32)
        for y in range(h):
            cur = px[y].astype(np.int32)
            ft = int(rng.integers(0, 5))
            left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]]) if rowb > bpp else np.zeros(rowb, np.int32)
            ul = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]]) if rowb > bpp else np.zeros(rowb, np.int32)
            if ft == 0:
                pred = 0
            elif ft == 1:
                pred = left
            elif ft == 2:
                pred = prev
            elif ft == 3:
                pred = (left + prev) // 2
            else:
                p = left + prev - ul
                pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - ul)
                pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, ul))
            stream.append(ft)
            stream += ((cur - pred) & 255).astype(np.uint8).tobytes()
            prev = cur
        ct = {1: 0, 2: 4, 3: 2, 4: 6}[ch]
        got = ctx.unfilter(w, h, bd, ct, 0, bytes(stream))
        assert np.array_equal(got.reshape(h, rowb), px), (w, h, ch, bd)


def test_unfilter_errors(ctx):
    raw = bytes([9]) + bytes(4) + bytes([0]) + bytes(4)  # 2 rows of 4 gray px, first filter byte invalid
    with pytest.raises(nb.NormError) as e:
        ctx.unfilter(4, 2, 8, 0, 0, raw)
    assert e.value.status == nb._lib.NORM_E_INVALID_ARG and "filter" in str(e.value)
    with pytest.raises(nb.NormError):  # stream length does not match the geometry
        ctx.unfilter(4, 2, 8, 0, 0, raw[:-1])
    with pytest.raises(nb.NormError):  # bad depth for RGB
        ctx.unfilter(4, 2, 4, 2, 0, raw)


# ---- batches: n streams of identical geometry, one CTA per (image, pass) ---------------------

def _groups():
    g = {}
    for name in VALID:
        w, h, bd, ct, il, raw, plte, trns = inflate_png(os.path.join(PNGSUITE, name + ".png"))
        if ct != 3 and len(trns) == 0:  # palette / colour key are per file; a batch shares them
            g.setdefault((w, h, bd, ct, il), []).append(name)
    return {k: v for k, v in g.items() if len(v) >= 2}


GROUPS = _groups()


@pytest.mark.parametrize("key", sorted(GROUPS), ids=lambda k: "%dx%d_bd%d_ct%d_il%d" % k)
def test_unfilter_batch_pngsuite(ctx, oracle, key):
    w, h, bd, ct, il = key
    names = GROUPS[key]
    streams = [inflate_png(os.path.join(PNGSUITE, n + ".png"))[5] for n in names]
    out = ctx.image_create(w, h, len(names) + 1)
    ctx.unfilter_expand_batch(w, h, bd, ct, il, streams, nb.NORM_EXPAND_SPEC, out, 1)
    got = out.download()
    for i, n in enumerate(names):
        d = oracle.png_decode(open(os.path.join(PNGSUITE, n + ".png"), "rb").read())
        assert np.array_equal(got[1 + i], oracle.expand_rgba8(d, oracle.EXPAND_SPEC)), n
    out.destroy()


def test_unfilter_batch_round_trip_through_gpu_filter(ctx):
    """frames -> GPU filter selection (all five types appear) -> batched GPU unfilter + expansion."""
    from conftest import smooth_rgba
    rng = np.random.default_rng(5)
    w, h, n = 257, 600, 9
    frames = [(smooth_rgba(w, h, seed=s)[..., :3] + rng.integers(0, 3, size=(h, w, 3))).astype(np.uint8) for s in range(n)]
    streams, used = [], set()
    for f in frames:
        st, ft = ctx.filter_scanlines(w, h, 8, 2, f)
        streams.append(st.tobytes())
        used |= set(ft.tolist())
    assert len(used) >= 3
    out = ctx.image_create(w, h, n)
    ctx.unfilter_expand_batch(w, h, 8, 2, 0, streams, nb.NORM_EXPAND_COMPAT, out)
    got = out.download()
    for i, f in enumerate(frames):
        assert np.array_equal(got[i][..., :3], f) and (got[i][..., 3] == 0).all()
    out.destroy()


def test_unfilter_batch_errors(ctx):
    good = bytes([0]) + bytes(4) + bytes([2]) + bytes(4)
    bad = bytes([0]) + bytes(4) + bytes([7]) + bytes(4)
    out = ctx.image_create(4, 2, 3)
    with pytest.raises(nb.NormError) as e:  # invalid filter byte in the last image of the batch
        ctx.unfilter_expand_batch(4, 2, 8, 0, 0, [good, good, bad], nb.NORM_EXPAND_SPEC, out)
    assert "filter" in str(e.value)
    with pytest.raises(nb.NormError):  # more streams than frames
        ctx.unfilter_expand_batch(4, 2, 8, 0, 0, [good] * 4, nb.NORM_EXPAND_SPEC, out)
    with pytest.raises(nb.NormError):  # one stream of the wrong length
        ctx.unfilter_expand_batch(4, 2, 8, 0, 0, [good, good[:-1]], nb.NORM_EXPAND_SPEC, out)
    ctx.unfilter_expand_batch(4, 2, 8, 0, 0, [good, good, good], nb.NORM_EXPAND_SPEC, out)
    assert (out.download()[..., :3] == 0).all()
    out.destroy()
