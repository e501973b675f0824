"""Pins the CPU oracle (oracle/norm_oracle.c) before anything is checked against it.

What pins what:
  * the reference's 13 golden blobs (reference tests src/png/decoder.rs:295-365) pin the decode
    front-end byte-for-byte: IHDR fields, gamma, pixel_buffer;
  * the same tests' rgb8() == image-crate to_rgb8() check is reproduced against PIL;
  * src/png/interlace.rs:89-124 pins the Adam7 pass sizes;
  * PIL / cv2 cross-check spec-mode expansion on all 161 valid PngSuite files;
  * the S-form (what CUDA must equal bit-exactly) stays within 1 LSB of the literal R-form.
Effects / SSIM have no reference fixture ("parity unpinned"): the checks here are internal
consistency against independent numpy evaluations of the cited formulas.
"""
import glob
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN, PNGSUITE, smooth_rgba, splitmix_rgba

BLOBS = ["basn0g08", "basn4a08", "basn6a08", "f00n0g08", "f00n2c08", "f01n0g08", "f01n2c08",
         "f02n0g08", "f02n2c08", "f03n0g08", "f03n2c08", "f04n0g08", "f04n2c08"]

ALL_PNGS = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(PNGSUITE, "*.png")))
CORRUPT = [n for n in ALL_PNGS if n.startswith("x")]
VALID = [n for n in ALL_PNGS if not n.startswith("x")]


def read_png(name):
    with open(os.path.join(PNGSUITE, name + ".png"), "rb") as f:
        return f.read()


def test_fixture_inventory():
    assert len(ALL_PNGS) == 175 and len(CORRUPT) == 14 and len(VALID) == 161


@pytest.mark.parametrize("name", BLOBS)
def test_golden_blob(oracle, name):
    """grammar.rs:203-218 blob layout: w,h u32be; depth, colour type, compression, filter,
    interlace u8; gamma u32be; pixel_buffer."""
    blob = open(os.path.join(PNGSUITE, name + ".blob"), "rb").read()
    w, h, bd, ct, comp, filt, inter, gamma = struct.unpack(">IIBBBBBI", blob[:17])
    png = oracle.png_decode(read_png(name))
    assert (png.width, png.height, png.bit_depth, png.color_type) == (w, h, bd, ct)
    assert (png.compression, png.filter, png.interlace, png.gamma) == (comp, filt, inter, gamma)
    assert png.pixels.tobytes() == blob[17:]


@pytest.mark.parametrize("name", BLOBS)
def test_rgb8_matches_library_decode(oracle, name):
    """decoder.rs:267-278: rgb8() equals an independent decoder's RGB8 (PIL stands in for the
    `image` crate; identical for 8-bit files)."""
    from PIL import Image
    png = oracle.png_decode(read_png(name))
    ref = np.asarray(Image.open(os.path.join(PNGSUITE, name + ".png")).convert("RGB"))
    assert np.array_equal(oracle.rgb8(png), ref)


def test_adam7_pass_sizes(oracle):
    """interlace.rs:89-124"""
    def counts(w, h):
        pw, ph = oracle.adam7_pass_sizes(w, h)
        return [a * b for a, b in zip(pw, ph)]
    assert counts(8, 8) == [1, 1, 2, 4, 8, 16, 32]
    assert counts(9, 9) == [4, 2, 3, 6, 10, 20, 36]
    assert counts(9, 10) == [4, 2, 3, 6, 10, 20, 45]
    # sizes the reference underflows on (interlace.rs:47): every pixel lands in exactly one pass
    for w in range(1, 12):
        for h in range(1, 12):
            assert sum(counts(w, h)) == w * h


@pytest.mark.parametrize("name", CORRUPT)
def test_corrupt_rejected(oracle, name):
    with pytest.raises(ValueError):
        oracle.png_decode(read_png(name))


def _pil_rgba(name):
    from PIL import Image
    im = Image.open(os.path.join(PNGSUITE, name + ".png"))
    im.load()
    return im


@pytest.mark.parametrize("name", VALID)
def test_spec_expansion_vs_independent_decoders(oracle, name):
    import cv2
    png = oracle.png_decode(read_png(name))
    got = oracle.expand_rgba8(png, oracle.EXPAND_SPEC)
    assert got.shape == (png.height, png.width, 4)
    path = os.path.join(PNGSUITE, name + ".png")
    if png.bit_depth == 16:
        # raw u16 from OpenCV, then the (v+128)/257 convention applied here
        raw = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        assert raw is not None and raw.dtype == np.uint16
        raw = raw.astype(np.uint32)
        d8 = ((raw + 128) // 257).astype(np.uint8)
        if d8.ndim == 2:
            exp_rgb = np.stack([d8] * 3, -1)
            exp_a = None
        elif d8.shape[2] == 3:
            exp_rgb, exp_a = d8[..., ::-1], None
        else:
            exp_rgb, exp_a = d8[..., [2, 1, 0]], d8[..., 3]
        assert np.array_equal(got[..., :3], exp_rgb)
        if exp_a is not None and png.color_type in (4, 6):
            assert np.array_equal(got[..., 3], exp_a)
        elif png.trns.size == 0:
            assert (got[..., 3] == 255).all()
        else:
            # colour key compared at 16 bits
            key = png.trns.reshape(-1, 2).astype(np.uint32)
            key = key[:, 0] * 256 + key[:, 1]
            r16 = raw if raw.ndim == 3 else raw[..., None]
            r16 = r16[..., ::-1] if r16.shape[2] == 3 else r16
            hit = np.all(r16[..., : key.size] == key, axis=-1)
            assert np.array_equal(got[..., 3], np.where(hit, 0, 255))
        return
    im = _pil_rgba(name)
    ref = np.asarray(im.convert("RGBA"))
    assert np.array_equal(got[..., :3], ref[..., :3]), name
    if png.color_type in (3, 4, 6) or png.trns.size == 0:
        assert np.array_equal(got[..., 3], ref[..., 3]), name
    else:
        # gray / RGB colour key: evaluate the PNG-spec rule independently at native depth
        if png.color_type == 0:
            bd = png.bit_depth
            key = ((int(png.trns[0]) << 8) | int(png.trns[1])) & ((1 << bd) - 1)
            scale = {1: 255, 2: 85, 4: 17, 8: 1}[bd]
            exp_a = np.where(got[..., 0] == key * scale, 0, 255)
        else:
            key = png.trns.reshape(3, 2).astype(np.int32)
            key = (key[:, 0] * 256 + key[:, 1]) & 0xFF
            exp_a = np.where(np.all(got[..., :3] == key, axis=-1), 0, 255)
        assert np.array_equal(got[..., 3], exp_a), name


def test_compat_expansion_semantics(oracle):
    """grammar.rs:123-155: alpha 0 for opaque types; only 8-bit ct 0/2/4/6."""
    n_ok = 0
    for name in VALID:
        png = oracle.png_decode(read_png(name))
        eligible = png.bit_depth == 8 and png.color_type != 3
        if not eligible:
            with pytest.raises(ValueError):
                oracle.expand_rgba8(png, oracle.EXPAND_COMPAT)
            continue
        n_ok += 1
        got = oracle.expand_rgba8(png, oracle.EXPAND_COMPAT)
        rgb = oracle.rgb8(png)
        assert np.array_equal(got[..., :3], rgb)
        if png.color_type in (0, 2):
            assert (got[..., 3] == 0).all()
        elif png.color_type == 4:
            assert np.array_equal(got[..., 3], png.pixels.reshape(png.height, png.width, 2)[..., 1])
        else:
            assert np.array_equal(got.reshape(-1), png.pixels)
    assert n_ok >= 40  # the 40 files the reference expands + interlaced/PLTE-carrying 8-bit ones


def test_obama_decodes(oracle):
    data = open(os.path.join(GOLDEN, "images", "obama.png"), "rb").read()
    png = oracle.png_decode(data)
    assert (png.width, png.height, png.bit_depth, png.color_type, png.gamma) == (1024, 683, 8, 2, 0)
    from PIL import Image
    ref = np.asarray(Image.open(os.path.join(GOLDEN, "images", "obama.png")).convert("RGB"))
    assert np.array_equal(oracle.rgb8(png), ref)


# ---------------------------------------------------------------- effects: S-form vs R-form

def _imgs():
    return [splitmix_rgba(61, 47), smooth_rgba(96, 70, seed=3), smooth_rgba(33, 129, seed=5)]


def test_front_lut_properties(oracle):
    ident = oracle.front_lut(oracle.CS_RAW, 0)
    assert np.array_equal(ident, np.arange(256, dtype=np.uint8))
    lin = oracle.front_lut(oracle.CS_REFERENCE, 0).astype(int)
    assert lin[0] == 0 and lin[255] == 255 and (np.diff(lin) >= 0).all()
    c = np.arange(256) / 255.0
    eotf = np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)
    assert np.abs(lin - eotf * 255).max() <= 0.5 + 1e-4
    g = oracle.front_lut(oracle.CS_RAW, 45455).astype(int)
    assert np.abs(g - 255 * c ** (1 / 0.45455)).max() <= 0.5 + 1e-3


@pytest.mark.parametrize("gray,inv,gamma", [(0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 1, 45455), (1, 0, 100000)])
def test_front_s_vs_r(oracle, gray, inv, gamma):
    for img in _imgs():
        for cs in (oracle.CS_REFERENCE, oracle.CS_RAW):
            s = oracle.front_s(img, cs, gamma, gray, inv).astype(int)
            r = oracle.front_r(img, cs, gamma, gray, inv).astype(int)
            assert np.abs(s - r).max() <= 1
            assert np.array_equal(s[..., 3], img[..., 3])


@pytest.mark.parametrize("radius", [1, 2, 3, 4, 7, 21, 39])
def test_blur_s_vs_r(oracle, radius):
    for img in _imgs():
        s = oracle.blur_s(img, radius).astype(int)
        r = oracle.blur_r(img, radius).astype(int)
        assert np.abs(s - r).max() <= 1, radius
    if radius == 1:
        assert np.array_equal(oracle.blur_s(img, 1), img)


def test_blur_taps(oracle):
    for r in (3, 7, 21, 39):
        taps, dmin = oracle.blur_taps(r)
        m = (r - 1) // 2
        assert dmin == -m and abs(float(taps.astype(np.float64).sum()) - 1) < 1e-6
        # kernel is centred at +0.5 px: w(d) == w(1-d)
        d = np.arange(dmin, dmin + r)
        for k, dk in enumerate(d):
            if dmin <= 1 - dk <= dmin + r - 1:
                assert taps[k] == taps[(1 - dk) - dmin]
    taps, dmin = oracle.blur_taps(21)
    # SURVEY appendix A
    assert abs(taps[0] - 0.01078) < 1e-5 and abs(taps[10] - 0.07931) < 1e-5 and abs(taps[11] - 0.07931) < 1e-5
    taps, dmin = oracle.blur_taps(4)
    assert dmin == -2 and taps[1] == taps[3]


def test_blur_constant_image_is_fixed_point(oracle):
    img = np.full((40, 50, 4), 0, np.uint8)
    img[...] = (13, 200, 255, 77)
    for r in (3, 21, 39):
        assert np.array_equal(oracle.blur_s(img, r), img)
        assert np.array_equal(oracle.blur_r(img, r), img)


def test_blur_numpy_restatement(oracle):
    """Independent numpy evaluation of the S-form definition (float64 accumulate is within
    rounding distance; compare within 1 LSB and mostly equal)."""
    img = smooth_rgba(50, 40, seed=9)
    r = 7
    taps, dmin = oracle.blur_taps(r)
    h, w, _ = img.shape
    xi = np.clip(np.arange(w)[:, None] + dmin + np.arange(r)[None, :], 0, w - 1)
    yi = np.clip(np.arange(h)[:, None] + dmin + np.arange(r)[None, :], 0, h - 1)
    H = (img.astype(np.float64)[:, xi, :] * taps.astype(np.float64)[None, None, :, None]).sum(2)
    V = (H[yi, :, :] * taps.astype(np.float64)[None, :, None, None]).sum(1)
    ref = np.clip(np.rint(V), 0, 255).astype(int)
    got = oracle.blur_s(img, r).astype(int)
    assert np.abs(ref - got).max() <= 1 and (ref != got).mean() < 0.01


def test_sobel_s_vs_r_and_numpy(oracle):
    for img in _imgs():
        s = oracle.sobel_s(img)
        r = oracle.sobel_r(img)
        assert np.abs(s.astype(int) - r.astype(int)).max() <= 1
        assert (s[..., 3] == 255).all() and np.array_equal(s[..., 0], s[..., 1]) and np.array_equal(s[..., 0], s[..., 2])
        # independent numpy evaluation incl. the reference's `-left` quirk (image_shader.wgsl:151)
        I = np.sqrt((img[..., :3].astype(np.float64) ** 2).sum(-1))
        P = np.pad(I, 1, mode="edge")
        h, w = I.shape
        def at(dx, dy):
            return P[1 + dy:1 + dy + h, 1 + dx:1 + dx + w]
        tl, t, tr = at(-1, 1), at(0, 1), at(1, 1)
        l, rr = at(-1, 0), at(1, 0)
        bl, b, br = at(-1, -1), at(0, -1), at(1, -1)
        gx = tl + 2 * l + bl - tr - 2 * rr - br
        gy = -l - 2 * t - tr + bl + 2 * b + br
        ref = np.clip(np.rint(np.minimum(np.sqrt(gx * gx + gy * gy), 255)), 0, 255).astype(int)
        assert np.abs(ref - s[..., 0].astype(int)).max() <= 1


@pytest.mark.parametrize("factor", [1, 16, 40])
def test_sharpen_s_equals_r(oracle, factor):
    for img in _imgs():
        assert np.array_equal(oracle.sharpen_s(img, factor), oracle.sharpen_r(img, factor))


def test_ssim_against_numpy(oracle):
    a = smooth_rgba(120, 90, seed=11)
    b = oracle.blur_s(a, 7)
    def luma(x):
        x = x.astype(np.float32)
        return (x[..., 0] * np.float32(0.29891) + x[..., 1] * np.float32(0.58661)) + x[..., 2] * np.float32(0.11448)
    x, y = luma(b).astype(np.float64).ravel(), luma(a).astype(np.float64).ravel()
    n = x.size
    mx, my = x.mean(), y.mean()
    vx, vy = ((x - mx) ** 2).sum() / (n - 1), ((y - my) ** 2).sum() / (n - 1)
    cov = ((x - mx) * (y - my)).sum() / (n - 1)
    C1, C2 = (0.01 * 255) ** 2, (0.03 * 255) ** 2
    C3 = C2 / 2
    sx, sy = np.sqrt(vx), np.sqrt(vy)
    ref = (2 * mx * my + C1) / (mx * mx + my * my + C1) * (2 * sx * sy + C2) / (vx + vy + C2) * (cov + C3) / (sx * sy + C3)
    got = oracle.ssim(b, a)
    assert abs(got - ref) < 1e-10
    m = oracle.ssim_moments(b, a)
    assert np.allclose(m, [x.sum(), y.sum(), (x * x).sum(), (y * y).sum(), (x * y).sum()], rtol=1e-12)
    assert abs(oracle.ssim_from_moments(m, n) - got) < 1e-12
    assert abs(oracle.ssim(a, a) - 1.0) < 1e-12
    # the reference's literal sequential-f32 arithmetic agrees to f32 noise at this size
    assert abs(oracle.ssim_f32_literal(b, a) - got) < 1e-4


def test_effects_priority_and_chain(oracle):
    img = smooth_rgba(64, 48, seed=21)
    h, w, _ = img.shape
    f = oracle.make_uniform(w, h, grayscale=1, blur=1, blur_radius=7, edge_detect=1, sharpen=1)
    A = oracle.front_s(img, oracle.CS_REFERENCE, 0, 1, 0)
    # image_shader.wgsl:164-174: blur wins over sharpen over edge, each from texture A
    assert np.array_equal(oracle.effects(img, f), oracle.blur_s(A, 7))
    f.blur = 0
    assert np.array_equal(oracle.effects(img, f), oracle.sharpen_s(A, 16))
    f.sharpen = 0
    assert np.array_equal(oracle.effects(img, f), oracle.sobel_s(A))
    f.edge_detect = 0
    assert np.array_equal(oracle.effects(img, f), A)
    f = oracle.make_uniform(w, h, grayscale=1, blur=1, blur_radius=7, edge_detect=1)
    out, s = oracle.chain(img, f)
    exp = oracle.sobel_s(oracle.blur_s(A, 7))
    assert np.array_equal(out, exp) and abs(s - oracle.ssim(exp, img)) < 1e-12
    outs, ss = oracle.chain_batch(np.stack([img, img[::-1].copy()]), f, oracle.CS_REFERENCE)
    assert np.array_equal(outs[0], out) and abs(ss[0] - s) < 1e-12


# ---------------------------------------------------------------- encoder-side filter selection

def _png_from_stream(w, h, bd, ct, stream):
    """Assemble a PNG around a filtered scanline stream (chunk.rs:17-96 layout: IHDR, one IDAT, IEND)."""
    import zlib

    def chunk(typ, body):
        return struct.pack(">I", len(body)) + typ + body + struct.pack(">I", zlib.crc32(typ + body) & 0xFFFFFFFF)
    ihdr = struct.pack(">IIBBBBB", w, h, bd, ct, 0, 0, 0)
    return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", ihdr) + chunk(b"IDAT", zlib.compress(bytes(stream), 1)) + chunk(b"IEND", b"")


def test_filter_selection_round_trips(oracle):
    """scanline_writer.rs:22-104 restated: the chosen filters decode back to the same pixels, both through
    the oracle's own decoder and through PIL (spec Paeth)."""
    import io
    from PIL import Image
    rng = np.random.default_rng(5)
    for (w, h, ch, bd, ct) in ((61, 40, 4, 8, 6), (33, 17, 3, 8, 2), (50, 9, 1, 8, 0), (20, 11, 2, 16, 4)):
        bpp, rowb = max(1, ch * bd // 8), (w * ch * bd + 7) // 8
        base = smooth_rgba(rowb, h, seed=w)[..., 0]  # smooth bytes so that all five filters get chosen somewhere
        px = (base + rng.integers(0, 3, size=base.shape)).astype(np.uint8)
        stream, ft = oracle.filter_scanlines(px, h, rowb, bpp, 0)
        d = oracle.png_decode(_png_from_stream(w, h, bd, ct, stream))
        assert np.array_equal(d.pixels, px.reshape(-1))
        if bd == 8:
            im = np.asarray(Image.open(io.BytesIO(_png_from_stream(w, h, bd, ct, stream))))
            assert np.array_equal(im.reshape(h, rowb), px)
        # the winner really has the smallest byte sum (first minimum)
        for y in range(h):
            row = stream[y * (rowb + 1) + 1:(y + 1) * (rowb + 1)]
            assert stream[y * (rowb + 1)] == ft[y] <= 4
            if ft[y] == 0:
                assert np.array_equal(row, px[y])


def test_reference_encoder_round_trip_on_obama(oracle):
    """encoder.rs:64-79 (`test_encode_round_trip`): decode(encode(decode(obama))) == decode(obama) holds in
    the reference even with its `a + b + c` Paeth (scanline_writer.rs:8); the bug-for-bug mode of the oracle
    must therefore round-trip obama through the spec decoder too."""
    data = open(os.path.join(GOLDEN, "images", "obama.png"), "rb").read()
    d = oracle.png_decode(data)
    rowb = d.width * 3
    for compat in (1, 0):
        stream, ft = oracle.filter_scanlines(d.pixels, d.height, rowb, 3, compat)
        back = oracle.png_decode(_png_from_stream(d.width, d.height, 8, 2, stream))
        assert np.array_equal(back.pixels, d.pixels), compat


# ---- box blur and shape overlay -------------------------------------

def test_box_blur_s_vs_r_and_numpy(oracle):
    img = splitmix_rgba(53, 31, seed=5)
    s, r = oracle.box_blur_s(img), oracle.box_blur_r(img)
    assert np.abs(s.astype(int) - r.astype(int)).max() <= 1
    pad = np.pad(img.astype(np.int64), ((1, 1), (1, 1), (0, 0)), mode="edge")
    acc = sum(pad[1 + dy:1 + dy + 31, 1 + dx:1 + dx + 53] for dy in (-1, 0, 1) for dx in (-1, 0, 1))
    ref = np.rint(acc / 9.0)  # no sum/9 lands on .5, so the rounding mode does not matter
    assert np.abs(s.astype(int) - ref).max() <= 0
    flat = np.full((9, 9, 4), 77, np.uint8)
    assert (oracle.box_blur_s(flat) == flat).all()


def _overlay_numpy(oracle, img, circles, selected, crosshair):
    """Vectorised f32 restatement of shape_shader.wgsl:41-89 + image_shader.wgsl:176-193 (no fma)."""
    f = np.float32
    h, w = img.shape[:2]
    eotf, thresh = oracle.srgb_tables()
    px = (np.arange(w, dtype=f) + f(0.5))[None, :].repeat(h, 0)
    py = (np.arange(h, dtype=f) + f(0.5))[:, None].repeat(w, 1)
    col = np.zeros((h, w, 4), f)

    def smooth(e0, e1, x):
        t = np.clip((x - e0) / (e1 - e0), f(0), f(1))
        return t * t * (f(3) - f(2) * t)

    def mix(a, b, t):
        return a * (f(1) - t) + b * t

    for i, (cx, cy, r) in enumerate(circles):
        cxp, cyp, rp = f(cx) * f(w), f(cy) * f(h), f(r) * f(min(w, h))
        d = np.sqrt((px - cxp) ** 2 + (py - cyp) ** 2, dtype=f)
        inside = d <= rp
        a = np.where(inside, f(1) - smooth(rp - f(1), rp, d), f(0)).astype(f)
        for c, v in enumerate((f(0), f(1), f(1))):
            col[..., c] = np.where(inside, mix(col[..., c], v, a), col[..., c])
        col[..., 3] = np.where(inside, mix(col[..., 3], a, a), col[..., 3])
        if selected == i:
            od = np.abs(d - rp)
            near = od <= f(3)
            a = np.where(near, f(1) - smooth(f(0), f(3), od), f(0)).astype(f)
            for c, v in enumerate((f(1), f(1), f(0))):
                col[..., c] = np.where(near, mix(col[..., c], v, a), col[..., c])
            col[..., 3] = np.where(near, mix(col[..., 3], a, a), col[..., 3])
    ca = np.rint(np.clip(col[..., 3], 0, 1) * f(255)).astype(np.int64)
    code = np.searchsorted(thresh[1:], (col[..., :3] * col[..., 3:4]).astype(f), side="right")
    sh = eotf[code]
    p = img.astype(f) / f(255)
    sa = (ca.astype(f) / f(255))[..., None]
    blended = np.concatenate([mix(p[..., :3], sh, sa), mix(p[..., 3:], sa, sa)], axis=2)
    p = np.where((ca > 0)[..., None], blended, p)
    hit = np.zeros((h, w), bool)
    if crosshair is not None:
        hit = np.sqrt((px - f(crosshair[0])) ** 2 + (py - f(crosshair[1])) ** 2, dtype=f) < f(crosshair[2])
        p[hit, 0], p[hit, 1], p[hit, 2] = 0, 1, 1
    out = np.rint(np.clip(p, 0, 1) * f(255)).astype(np.uint8)
    untouched = (ca == 0) & ~hit
    out[untouched] = img[untouched]
    return out


def test_overlay_against_numpy_and_properties(oracle):
    img = splitmix_rgba(160, 120, seed=9)
    circles = [(0.5, 0.5, 0.2), (0.3, 0.4, 0.1), (0.9, 0.95, 0.15), (0.55, 0.5, 0.05)]
    out = oracle.overlay(img, circles, selected=1, crosshair=(20.0, 100.0, 6.5))
    ref = _overlay_numpy(oracle, img, circles, 1, (20.0, 100.0, 6.5))
    d = np.abs(out.astype(int) - ref.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.01  # fma vs separate rounding only
    # interior of a filled circle is opaque cyan with alpha 255; far away nothing changes
    assert tuple(out[60, 80]) == (0, 255, 255, 255)
    assert (out[0:10, 0:10] == img[0:10, 0:10]).all()
    # crosshair disc sets rgb = (0, 1, 1) and keeps alpha
    assert tuple(out[100, 20][:3]) == (0, 255, 255) and out[100, 20][3] == img[100, 20][3]
    # the selected circle gets a yellow ring at its radius: r = 0.1 * 120 = 12 px around (48, 48)
    ring = out[48, 48 + 12]
    assert ring[0] > 200 and ring[1] > 200 and ring[2] < 60
    # no circles, no crosshair: identity
    assert (oracle.overlay(img) == img).all()
    # blending is in list order: the later small circle lies inside the first one, result unchanged cyan
    assert tuple(out[60, 88]) == (0, 255, 255, 255)


def test_overlay_rejects_bad_arguments(oracle):
    img = splitmix_rgba(8, 8, seed=1)
    with pytest.raises(ValueError):
        oracle.overlay(img, [(0.5, 0.5, 0.1)] * 257)


def test_srgb_tables(oracle):
    eotf, thresh = oracle.srgb_tables()
    assert eotf[0] == 0 and eotf[255] == 1 and (np.diff(eotf) > 0).all() and (np.diff(thresh[1:]) > 0).all()
    # every code decodes to a value that encodes back to itself
    code = np.searchsorted(thresh[1:], eotf, side="right")
    assert (code == np.arange(256)).all()


def test_srgb_surface_encode_table(oracle):
    """gpu_state.rs:94-99: the surface encode of a stored linear byte.  The threshold table must agree with the direct
    formula round(255 * OETF(c / 255)) for every byte, invert the load-time decode (texture.rs:36) on every code that
    survives the 8-bit linear store, and be monotone."""
    lut = oracle.srgb_encode_lut()
    x = np.arange(256) / 255.0
    direct = np.rint(255 * np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(x, 1 / 2.4) - 0.055)).astype(int)
    assert np.array_equal(lut.astype(int), direct)
    assert (np.diff(lut.astype(int)) >= 0).all() and lut[0] == 0 and lut[255] == 255
    dec = oracle.front_lut(oracle.CS_REFERENCE, 0)  # sRGB code -> stored linear byte
    hi = np.arange(64, 256)  # above the dark end the 8-bit linear store keeps codes apart well enough to come back within 1
    assert np.abs(lut[dec[hi]].astype(int) - hi).max() <= 3
    img = np.arange(4 * 256, dtype=np.uint8).reshape(1, 256, 4)
    enc = oracle.encode_srgb(img)
    assert np.array_equal(enc[..., 3], img[..., 3]) and np.array_equal(enc[..., :3], lut[img[..., :3]])
