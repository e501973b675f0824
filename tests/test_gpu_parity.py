"""GPU parity: every CUDA stage, called through the C ABI, against the CPU oracle on the same
inputs.  Bars (BASELINE.json north_star): expansion bit-exact; filters bit-exact against
the S-form oracle (hence within 1 LSB of the literal shader order, tests/test_oracle.py);
SSIM within 1e-5 absolute (we hold 1e-9).
"""
import glob
import os

import ctypes as C

import numpy as np
import pytest

import norm_b200 as nb
from conftest import GOLDEN, PNGSUITE, smooth_rgba, splitmix_rgba

pytestmark = pytest.mark.gpu

SSIM_TOL = 1e-5     # the stated tolerance
SSIM_TIGHT = 1e-9   # what identical f32 lumas + f64 moments actually give


@pytest.fixture(scope="module")
def ctx():
    c = nb.Context(0)
    yield c
    c.close()


def to_png(d):
    return nb.Png(d.width, d.height, d.bit_depth, d.color_type, d.pixels, d.gamma, d.plte, d.trns)


def read_png(oracle, path):
    with open(path, "rb") as f:
        return oracle.png_decode(f.read())


VALID = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(PNGSUITE, "*.png"))
               if not os.path.basename(p).startswith("x"))

# odd sizes exercise ragged rows (w % 4 != 0), tiles smaller / larger than the CTA tile
SIZES = [(1, 1), (3, 2), (5, 7), (33, 17), (64, 32), (65, 33), (130, 70), (257, 129), (200, 331)]


def run_stage(ctx, fn, img, *args):
    src = ctx.image_from_host(img)
    dst = ctx.image_create(src.width, src.height, src.frames)
    try:
        fn(*args, src, dst)
        return dst.download()
    finally:
        src.destroy(), dst.destroy()


# ------------------------------------------------------------------ expansion (config 2)

@pytest.mark.parametrize("name", VALID)
def test_expand_spec_pngsuite(ctx, oracle, name):
    d = read_png(oracle, os.path.join(PNGSUITE, name + ".png"))
    got = to_png(d).rgba8(ctx, nb.NORM_EXPAND_SPEC)
    assert np.array_equal(got, oracle.expand_rgba8(d, oracle.EXPAND_SPEC))


def test_expand_compat_pngsuite(ctx, oracle):
    n = 0
    for name in VALID:
        d = read_png(oracle, os.path.join(PNGSUITE, name + ".png"))
        if d.bit_depth == 8 and d.color_type != 3:
            got = to_png(d).rgba8(ctx, nb.NORM_EXPAND_COMPAT)
            assert np.array_equal(got, oracle.expand_rgba8(d, oracle.EXPAND_COMPAT)), name
            n += 1
        else:
            with pytest.raises(nb.NormError) as e:
                to_png(d).rgba8(ctx, nb.NORM_EXPAND_COMPAT)
            assert e.value.status == nb._lib.NORM_E_UNSUPPORTED
    assert n >= 40


def test_expand_synthetic_all_types(ctx, oracle):
    """Every (colour type, depth) at ragged widths incl. tRNS keys and short palettes."""
    rng = np.random.default_rng(7)
    for ct, depths in ((0, (1, 2, 4, 8, 16)), (2, (8, 16)), (3, (1, 2, 4, 8)), (4, (8, 16)), (6, (8, 16))):
        for bd in depths:
            for (w, h) in ((1, 1), (7, 3), (37, 5), (130, 9)):
                ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ct]
                rowb = (w * ch * bd + 7) // 8
                pix = rng.integers(0, 256, size=rowb * h, dtype=np.uint8)
                plte = trns = None
                if ct == 3:
                    n = min(1 << bd, 200)
                    plte = rng.integers(0, 256, size=3 * n, dtype=np.uint8)
                    trns = rng.integers(0, 256, size=n // 2, dtype=np.uint8)
                elif ct == 0:
                    trns = np.array([0, 1 if bd < 16 else 7], np.uint8)
                elif ct == 2:
                    trns = np.array([0, pix[0], 0, pix[min(1, pix.size - 1)], 0, pix[min(2, pix.size - 1)]], np.uint8)

                class D:
                    pass
                d = D()
                d.width, d.height, d.bit_depth, d.color_type, d.gamma = w, h, bd, ct, 0
                d.compression = d.filter = d.interlace = 0
                d.pixels, d.plte = pix, plte if plte is not None else np.zeros(0, np.uint8)
                d.trns = trns if trns is not None else np.zeros(0, np.uint8)
                od = oracle.DecodedPng.__new__(oracle.DecodedPng)
                od.__dict__.update(d.__dict__)
                exp = oracle.expand_rgba8(od, oracle.EXPAND_SPEC)
                got = to_png(od).rgba8(ctx, nb.NORM_EXPAND_SPEC)
                assert np.array_equal(got, exp), (ct, bd, w, h)


def test_expand_errors(ctx):
    pix = np.zeros(16, np.uint8)
    with pytest.raises(nb.NormError):  # bad depth for RGB
        nb.Png(2, 2, 4, 2, pix).rgba8(ctx, nb.NORM_EXPAND_SPEC)
    with pytest.raises(nb.NormError):  # short buffer
        nb.Png(4, 4, 8, 6, pix).rgba8(ctx, nb.NORM_EXPAND_SPEC)
    with pytest.raises(nb.NormError):  # palette without PLTE
        nb.Png(4, 4, 8, 3, pix).rgba8(ctx, nb.NORM_EXPAND_SPEC)


# ------------------------------------------------------------------ single stages

@pytest.mark.parametrize("gray,inv,gamma,cs", [(0, 0, 0, 0), (1, 0, 0, 0), (0, 1, 0, 1), (1, 1, 45455, 0),
                                               (1, 0, 0, 1), (0, 0, 250000, 1), (1, 1, 100000, 1)])
def test_front(ctx, oracle, gray, inv, gamma, cs):
    for (w, h) in SIZES:
        img = splitmix_rgba(w, h, seed=w * 131 + h)
        f = nb.FeatureUniform(w, h, gamma)
        f.raw.grayscale, f.raw.invert = gray, inv
        got = run_stage(ctx, lambda s, d: ctx.front(f, s, d, cs), img)[0]
        assert np.array_equal(got, oracle.front_s(img, cs, gamma, gray, inv)), (w, h)


@pytest.mark.parametrize("radius", [1, 2, 3, 4, 5, 7, 9, 21, 39, 40, 63])
def test_blur(ctx, oracle, radius):
    sizes = SIZES if radius in (3, 21) else [(1, 1), (5, 7), (65, 33), (200, 131)]
    for (w, h) in sizes:
        img = smooth_rgba(w, h, seed=radius + w) if (w + h) % 2 else splitmix_rgba(w, h, seed=radius)
        got = run_stage(ctx, lambda s, d: ctx.blur(radius, s, d), img)[0]
        exp = oracle.blur_s(img, radius)
        assert np.array_equal(got, exp), (radius, w, h, int(np.abs(got.astype(int) - exp).max()))


def test_blur_bad_radius(ctx):
    src = ctx.image_from_host(splitmix_rgba(8, 8))
    dst = ctx.image_create(8, 8, 1)
    for r in (0, 64, 1000):
        with pytest.raises(nb.NormError) as e:
            ctx.blur(r, src, dst)
        assert e.value.status == nb._lib.NORM_E_INVALID_ARG
    with pytest.raises(nb.NormError):  # aliasing
        ctx.blur(3, src, src)
    src.destroy(), dst.destroy()


def test_sobel(ctx, oracle):
    for (w, h) in SIZES:
        for img in (splitmix_rgba(w, h, seed=5), smooth_rgba(w, h, seed=6)):
            got = run_stage(ctx, ctx.sobel, img)[0]
            assert np.array_equal(got, oracle.sobel_s(img)), (w, h)


@pytest.mark.parametrize("factor", [1, 16, 40])
def test_sharpen(ctx, oracle, factor):
    for (w, h) in SIZES:
        img = smooth_rgba(w, h, seed=factor)
        got = run_stage(ctx, lambda s, d: ctx.sharpen(factor, s, d), img)[0]
        assert np.array_equal(got, oracle.sharpen_s(img, factor)), (w, h)


def test_ssim(ctx, oracle):
    for (w, h) in SIZES[1:] + [(1024, 683)]:
        a = smooth_rgba(w, h, seed=w)
        b = oracle.blur_s(a, 7)
        da, db = ctx.image_from_host(b), ctx.image_from_host(a)
        got = ctx.ssim(da, db)[0]
        exp = oracle.ssim(b, a)
        assert abs(got - exp) <= SSIM_TIGHT and abs(got - exp) <= SSIM_TOL, (w, h, got, exp)
        m = ctx.ssim_moments(da, db)[0]
        assert np.allclose(m, oracle.ssim_moments(b, a), rtol=1e-13, atol=0)
        assert abs(nb.ssim_from_moments(m, w * h) - exp) <= SSIM_TIGHT
        assert abs(ctx.ssim(da, da)[0] - 1.0) <= 1e-12
        da.destroy(), db.destroy()
    # 1x1: N-1 = 0 -> NaN, as the reference's arithmetic gives (ssim.rs:169,180)
    one = ctx.image_from_host(splitmix_rgba(1, 1))
    assert np.isnan(ctx.ssim(one, one)[0])
    # dimension mismatch (ssim.rs:92-95)
    other = ctx.image_from_host(splitmix_rgba(2, 1))
    with pytest.raises(nb.NormError) as e:
        ctx.ssim(one, other)
    assert e.value.status == nb._lib.NORM_E_DIM_MISMATCH
    one.destroy(), other.destroy()


def test_ssim_is_deterministic(ctx):
    a, b = ctx.image_from_host(splitmix_rgba(640, 360, seed=1)), ctx.image_from_host(splitmix_rgba(640, 360, seed=2))
    r = [ctx.ssim_moments(a, b).tobytes() for _ in range(3)]
    assert r[0] == r[1] == r[2]
    a.destroy(), b.destroy()


# ------------------------------------------------------------------ pipelines

def uniform(w, h, **kw):
    f = nb.FeatureUniform(w, h, kw.pop("gamma", 0))
    for k, v in kw.items():
        setattr(f.raw, k, v)
    return f


def oracle_uniform(oracle, f):
    r = f.raw
    return oracle.make_uniform(r.width, r.height, r.gamma, r.grayscale, r.invert, r.blur, r.blur_radius, r.sharpen,
                               r.sharpen_factor, r.edge_detect)


@pytest.mark.parametrize("kw", [
    dict(), dict(grayscale=1), dict(blur=1, blur_radius=7), dict(grayscale=1, blur=1, blur_radius=21),
    dict(edge_detect=1), dict(sharpen=1), dict(blur=1, sharpen=1, edge_detect=1, blur_radius=3),
    dict(sharpen=1, edge_detect=1, invert=1), dict(grayscale=2, blur=2, edge_detect=1, gamma=45455),
])
def test_effects_reference_semantics(ctx, oracle, kw):
    """norm_cuda_effects == one reference frame: priority blur > sharpen > edge; flags are on iff == 1."""
    for (w, h) in [(5, 7), (130, 70), (200, 131)]:
        img = smooth_rgba(w, h, seed=11)
        f = uniform(w, h, **kw)
        for cs in (nb.NORM_CS_REFERENCE, nb.NORM_CS_RAW):
            got = run_stage(ctx, lambda s, d: ctx.effects(f, s, d, cs), img)[0]
            assert np.array_equal(got, oracle.effects(img, oracle_uniform(oracle, f), cs)), (kw, w, h, cs)


def test_compute_sim_follows_the_colour_type(ctx, oracle):
    """Png::compute_sim (ssim.rs:91-106): the luma of each image is the branch of luma_buffer its colour type
    takes (ssim.rs:18-87) -- RGB weights, the gray byte, or gray byte / 255 for gray + alpha -- and the two images
    may differ in it.  Palette images raise, as the reference's `unimplemented!` does."""
    rng = np.random.default_rng(5)
    w, h = 97, 61
    mk = {0: lambda: rng.integers(0, 256, size=w * h, dtype=np.uint8), 2: lambda: rng.integers(0, 256, size=w * h * 3, dtype=np.uint8),
          4: lambda: rng.integers(0, 256, size=w * h * 2, dtype=np.uint8), 6: lambda: rng.integers(0, 256, size=w * h * 4, dtype=np.uint8)}
    mode_of = {0: 1, 2: 0, 4: 2, 6: 0}
    for ct_a in (0, 2, 4, 6):
        for ct_b in (0, 2, 4, 6):
            a, b = nb.Png(w, h, 8, ct_a, mk[ct_a]()), nb.Png(w, h, 8, ct_b, mk[ct_b]())
            got = a.compute_sim(b, ctx)
            ea, eb = a.rgba8(ctx), b.rgba8(ctx)
            assert abs(got - oracle.ssim_luma(ea, eb, mode_of[ct_a], mode_of[ct_b])) <= SSIM_TIGHT, (ct_a, ct_b)
    ga = nb.Png(w, h, 8, 4, mk[4]())
    assert abs(ga.compute_sim(ga, ctx) - 1.0) < 1e-9                        # identical images: 1 whatever the scale
    with pytest.raises(nb.NormError):
        nb.Png(4, 4, 8, 3, np.zeros(16, np.uint8), plte=np.zeros(6, np.uint8)).compute_sim(ga, ctx)
    with pytest.raises(nb.NormError):
        ga.compute_sim(nb.Png(w + 1, h, 8, 4, np.zeros((w + 1) * h * 2, np.uint8)), ctx)   # ensure! on the dimensions


def test_expand_device_resident_batch(ctx, oracle):
    """norm_cuda_expand_rgba8_device: scanlines already in device memory at a 16-byte pitch, several images per
    launch, every colour type / depth, ragged widths -- bit-exact against the oracle's expansion."""
    import torch
    rng = np.random.default_rng(77)
    ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}
    for (ct, bd) in [(0, 1), (0, 2), (0, 4), (0, 8), (0, 16), (2, 8), (2, 16), (3, 1), (3, 2), (3, 4), (3, 8), (4, 8), (4, 16), (6, 8), (6, 16)]:
        for (w, h) in [(1, 1), (5, 3), (33, 7), (130, 9)]:
            nf = 3
            rowb = (w * ch[ct] * bd + 7) // 8
            pitch = (rowb + 15) // 16 * 16
            fstride = pitch * h + 64
            host = rng.integers(0, 256, size=(nf, fstride), dtype=np.uint8)
            plte = rng.integers(0, 256, size=3 * (1 << min(bd, 8)), dtype=np.uint8) if ct == 3 else None
            trns = (rng.integers(0, 256, size=(1 << min(bd, 8)) // 2 + 1, dtype=np.uint8) if ct == 3 else
                    np.array([0, host[0, 0] >> (8 - bd) if bd < 8 else host[0, 0]], np.uint8) if ct == 0 and bd <= 8 else None)
            dev = torch.from_numpy(host).cuda()
            torch.cuda.synchronize()
            for mode in ([nb.NORM_EXPAND_SPEC] + ([nb.NORM_EXPAND_COMPAT] if bd == 8 and ct != 3 else [])):
                out = ctx.image_create(w, h, nf + 1)
                meta = nb.Png(w, h, bd, ct, np.zeros(1, np.uint8), plte=plte, trns=trns)
                ctx.expand_rgba8_device(meta, dev.data_ptr(), pitch, fstride, nf, mode, out, 1)
                ctx.sync()
                got = out.download()
                for i in range(nf):
                    tight = np.ascontiguousarray(host[i, :pitch * h].reshape(h, pitch)[:, :rowb]).reshape(-1)
                    d = oracle.DecodedPng.__new__(oracle.DecodedPng)
                    d.__dict__.update(width=w, height=h, bit_depth=bd, color_type=ct, compression=0, filter=0, interlace=0, gamma=0,
                                      pixels=tight, plte=plte if plte is not None else np.zeros(0, np.uint8),
                                      trns=trns if trns is not None else np.zeros(0, np.uint8))
                    exp = oracle.expand_rgba8(d, oracle.EXPAND_SPEC if mode == nb.NORM_EXPAND_SPEC else oracle.EXPAND_COMPAT)
                    assert np.array_equal(got[1 + i], exp), (ct, bd, w, h, mode, i)
                out.destroy()


def test_surface_srgb_encode_on_output(ctx, oracle):
    """NORM_OUT_SRGB: what a norm user sees is the fragment result sRGB-encoded by the surface
    (gpu_state.rs:94-99).  The library's table is the oracle's; effects, chain, host chain and a banded step with
    the flag equal the oracle's linear result put through it; alpha is untouched and the SSIM stays that of the
    linear result."""
    lut = (C.c_uint8 * 256)()
    assert nb.load().norm_srgb_encode_lut(lut) == 0
    assert np.array_equal(np.frombuffer(bytes(lut), np.uint8), oracle.srgb_encode_lut())
    for (w, h) in [(5, 7), (130, 70), (333, 200)]:
        img = smooth_rgba(w, h, seed=21)
        for kw in (dict(), dict(grayscale=1, blur=1, blur_radius=21), dict(edge_detect=1), dict(blur=1, blur_radius=7, edge_detect=1),
                   dict(blur=1, blur_radius=21, edge_detect=1)):
            f = uniform(w, h, **kw)
            for cs in (nb.NORM_CS_REFERENCE, nb.NORM_CS_RAW):
                got = run_stage(ctx, lambda s, d: ctx.effects(f, s, d, cs | nb.NORM_OUT_SRGB), img)[0]
                assert np.array_equal(got, oracle.encode_srgb(oracle.effects(img, oracle_uniform(oracle, f), cs))), (kw, w, h, cs)
                src, dst = ctx.image_from_host(img[None]), ctx.image_create(w, h, 1)
                s_lin = ctx.chain(f, src, dst, cs)
                s_enc = ctx.chain(f, src, dst, cs | nb.NORM_OUT_SRGB)
                eo, es = oracle.chain(img, oracle_uniform(oracle, f), cs)
                assert np.array_equal(dst.download()[0], oracle.encode_srgb(eo)), (kw, w, h, cs)
                assert s_lin[0] == s_enc[0] and abs(s_enc[0] - es) <= SSIM_TIGHT
                hin, hout = np.ascontiguousarray(img[None]), np.empty((1, h, w, 4), np.uint8)
                ctx.chain_host(f, hin, hout, 1, cs | nb.NORM_OUT_SRGB)
                assert np.array_equal(hout[0], oracle.encode_srgb(eo))
                src.destroy(), dst.destroy()
    # banded: the owned rows come out encoded, whichever path (windowed or whole band) ran
    w, h, nranks = 200, 150, 3
    img = smooth_rgba(w, h, seed=22)
    f = uniform(w, h, blur=1, blur_radius=21, edge_detect=1)
    ctxs = [nb.Context(0) for _ in range(nranks)]
    bands = [nb.Band(c, f, w, h, nranks, r) for r, c in enumerate(ctxs)]
    handles = [b.export() for b in bands]
    for b in bands:
        b.connect(handles)
        ext = np.zeros((b.src.height, w, 4), np.uint8)
        ext[b.halo_up:b.halo_up + b.rows] = img[b.row0:b.row0 + b.rows]
        b.src.upload(ext)
        b.publish()
    for b in bands:
        b.step_async(nb.NORM_CS_RAW | nb.NORM_OUT_SRGB)
    for b in bands:
        b.sync()
    got = np.concatenate([b.own_rows_of(b.dst) for b in bands])
    assert np.array_equal(got, oracle.encode_srgb(oracle.chain(img, oracle_uniform(oracle, f), oracle.CS_RAW)[0]))
    for b in bands:
        b.destroy()
    for c in ctxs:
        c.close()


CHAINS = [
    dict(blur=1, blur_radius=21, edge_detect=1),                  # BASELINE config 4 chain
    dict(grayscale=1, blur=1, blur_radius=21, edge_detect=1),     # config 3 chain
    dict(grayscale=1, blur=1, blur_radius=21),                    # config 1 effect
    dict(blur=1, blur_radius=3, edge_detect=1), dict(blur=1, blur_radius=7, edge_detect=1),
    dict(blur=1, blur_radius=39, edge_detect=1), dict(blur=1, blur_radius=5, edge_detect=1),
    dict(edge_detect=1), dict(blur=1, blur_radius=21), dict(), dict(invert=1, blur=1, sharpen=1, edge_detect=1, blur_radius=7),
    dict(grayscale=1, gamma=45455, blur=1, blur_radius=9, edge_detect=1),
]


@pytest.mark.parametrize("kw", CHAINS)
def test_chain(ctx, oracle, kw):
    for (w, h) in [(1, 1), (5, 7), (64, 32), (130, 70), (333, 200)]:
        frames = np.stack([smooth_rgba(w, h, seed=3), splitmix_rgba(w, h, seed=4, alpha=255),
                           smooth_rgba(w, h, seed=5)[::-1].copy()])
        f = uniform(w, h, **kw)
        for cs in (nb.NORM_CS_RAW, nb.NORM_CS_REFERENCE):
            src = ctx.image_from_host(frames)
            dst = ctx.image_create(w, h, 3)
            s = ctx.chain(f, src, dst, cs)
            got = dst.download()
            eo, es = oracle.chain_batch(frames, oracle_uniform(oracle, f), cs)
            assert np.array_equal(got, eo), (kw, w, h, cs, int(np.abs(got.astype(int) - eo).max()))
            if w * h > 1:
                assert np.abs(s - es).max() <= SSIM_TIGHT, (kw, w, h, s, es)
            src.destroy(), dst.destroy()


def test_chain_host_pipeline(ctx, oracle):
    """Host-buffer entry point (the e2e path): chunked upload / run / download."""
    w, h, n = 160, 90, 7
    frames = np.stack([smooth_rgba(w, h, seed=i) for i in range(n)])
    f = uniform(w, h, blur=1, blur_radius=21, edge_detect=1)
    out = np.empty_like(frames)
    s = ctx.chain_host(f, frames, out, n, nb.NORM_CS_RAW)
    eo, es = oracle.chain_batch(frames, oracle_uniform(oracle, f), oracle.CS_RAW)
    assert np.array_equal(out, eo) and np.abs(s - es).max() <= SSIM_TIGHT
    # pinned buffers
    pin, ph = ctx.host_alloc(frames.nbytes)
    pout, poh = ctx.host_alloc(frames.nbytes)
    pin[:] = frames.reshape(-1)
    s2 = ctx.chain_host(f, pin, pout, n, nb.NORM_CS_RAW)
    assert np.array_equal(pout.reshape(frames.shape), eo) and np.array_equal(s, s2)
    ctx.host_free(ph), ctx.host_free(poh)


def test_balanced_partition_of_one_frame(oracle, monkeypatch):
    """One giant frame is cut into one contiguous range of strip-rows per CTA (items that start and end anywhere in a
    strip, empty items, SSIM slots written as zero).  NORM_CUDA_BALANCED=2 forces that partition with ranges as short
    as 7 rows, so small images exercise every case; results must equal the oracle and the static partition."""
    monkeypatch.setenv("NORM_CUDA_BALANCED", "2")
    with nb.Context(0) as c:
        for (w, h) in [(1, 1), (5, 7), (130, 70), (333, 200), (1000, 517)]:
            img = smooth_rgba(w, h, seed=w + h)
            for kw in (dict(blur=1, blur_radius=21, edge_detect=1), dict(grayscale=1, blur=1, blur_radius=7, edge_detect=1),
                       dict(blur=1, blur_radius=3, edge_detect=1), dict(blur=1, blur_radius=21), dict(grayscale=1, blur=1, blur_radius=21)):
                f = uniform(w, h, **kw)
                src, dst = c.image_from_host(img[None]), c.image_create(w, h, 1)
                s = c.chain(f, src, dst, nb.NORM_CS_RAW)
                eo, es = oracle.chain(img, oracle_uniform(oracle, f), oracle.CS_RAW)
                assert np.array_equal(dst.download()[0], eo), (w, h, kw)
                assert abs(s[0] - es) <= SSIM_TIGHT or (w * h == 1), (w, h, kw)
                src.destroy(), dst.destroy()


def test_chain_r3_window_kernel_ragged(oracle, monkeypatch):
    """The chain at r = 3 runs chain_win3_kernel and the blur alone blur_win3_kernel (register windows, one warp per
    120-column tile): widths around the tile and chunk boundaries, heights around the tile-row counts and the three-step unroll, every front-stage
    combination, a batch -- against the oracle, and against the pipelined kernel (NORM_CUDA_WIN3=0)."""
    sizes = [(1, 1), (2, 3), (3, 2), (4, 1), (5, 19), (119, 18), (120, 20), (121, 37), (123, 74), (240, 5), (243, 145), (1000, 77)]
    kws = [dict(blur=1, blur_radius=3, edge_detect=1), dict(grayscale=1, blur=1, blur_radius=3, edge_detect=1),
           dict(invert=1, gamma=45455, blur=1, blur_radius=3, edge_detect=1),
           # the blur alone (blur_win3_kernel: all four channels; two after grayscale)
           dict(blur=1, blur_radius=3), dict(grayscale=1, blur=1, blur_radius=3), dict(invert=1, gamma=250000, blur=1, blur_radius=3)]
    results = {}
    for win3 in ("1", "0"):
        monkeypatch.setenv("NORM_CUDA_WIN3", win3)
        with nb.Context(0) as c:
            for (w, h) in sizes:
                img = smooth_rgba(w, h, seed=3 * w + h)
                for ki, kw in enumerate(kws):
                    f = uniform(w, h, **kw)
                    src, dst = c.image_from_host(img[None]), c.image_create(w, h, 1)
                    s = c.chain(f, src, dst, nb.NORM_CS_RAW)
                    got = dst.download()[0]
                    if win3 == "1":
                        eo, es = oracle.chain(img, oracle_uniform(oracle, f), oracle.CS_RAW)
                        assert np.array_equal(got, eo), (w, h, kw)
                        assert abs(s[0] - es) <= SSIM_TIGHT or (w * h == 1), (w, h, kw)
                        results[(w, h, ki)] = (got, s[0])
                    else:
                        assert np.array_equal(got, results[(w, h, ki)][0]), (w, h, kw)
                        assert abs(s[0] - results[(w, h, ki)][1]) <= SSIM_TIGHT or (w * h == 1)
                    src.destroy(), dst.destroy()
            if win3 == "1":  # a batch: tiles of several frames in one launch
                frames = np.stack([smooth_rgba(250, 91, seed=70 + i) for i in range(5)])
                f = uniform(250, 91, blur=1, blur_radius=3, edge_detect=1)
                eo, es = oracle.chain_batch(frames, oracle_uniform(oracle, f), oracle.CS_RAW)
                src, dst = c.image_from_host(frames), c.image_create(250, 91, 5)
                s = c.chain(f, src, dst, nb.NORM_CS_RAW)
                assert np.array_equal(dst.download(), eo) and np.abs(s - es).max() <= SSIM_TIGHT
                src.destroy(), dst.destroy()


def test_large_radius_configurations_ragged(oracle, monkeypatch):
    """The shift-form V pass and the one-CTA-per-SM configurations (RGB chain r >= 33, RGBA blur r >= 37; 34- / 38-row
    batches, 12- / 8-column H items) on sizes around their strip, batch and item boundaries, tiny images included, with
    the static partition and with the balanced partition of one frame forced."""
    sizes = [(1, 1), (7, 5), (119, 33), (121, 35), (130, 69), (241, 40), (500, 103)]
    kws = [dict(blur=1, blur_radius=23, edge_detect=1), dict(blur=1, blur_radius=31, edge_detect=1),
           dict(blur=1, blur_radius=33, edge_detect=1), dict(invert=1, blur=1, blur_radius=39, edge_detect=1),
           dict(grayscale=1, blur=1, blur_radius=39, edge_detect=1), dict(blur=1, blur_radius=29), dict(blur=1, blur_radius=37),
           dict(blur=1, blur_radius=39)]
    for balanced in ("1", "2"):
        monkeypatch.setenv("NORM_CUDA_BALANCED", balanced)
        with nb.Context(0) as c:
            for (w, h) in sizes:
                img = smooth_rgba(w, h, seed=5 * w + h)
                for kw in kws:
                    f = uniform(w, h, **kw)
                    src, dst = c.image_from_host(img[None]), c.image_create(w, h, 1)
                    s = c.chain(f, src, dst, nb.NORM_CS_RAW)
                    eo, es = oracle.chain(img, oracle_uniform(oracle, f), oracle.CS_RAW)
                    assert np.array_equal(dst.download()[0], eo), (balanced, w, h, kw)
                    assert abs(s[0] - es) <= SSIM_TIGHT or (w * h == 1), (balanced, w, h, kw)
                    src.destroy(), dst.destroy()


def test_multi_chunk_paths_bit_exact(oracle, monkeypatch):
    """The chunk loops only large batches reach: the host pipeline's slot reuse (chunk index mod 3) and the unfused
    chain's ping-pong temporaries rotating across chunks.  NORM_CUDA_CHUNK_BYTES (read when a context is created)
    shrinks the chunk to one frame, so 11 frames make 11 chunks; outputs and SSIM must still equal the oracle."""
    w, h, n = 160, 90, 11
    monkeypatch.setenv("NORM_CUDA_CHUNK_BYTES", str(160 * 4 * 90 + 256 * 90))  # one (pitched) frame per chunk
    frames = np.stack([smooth_rgba(w, h, seed=40 + i) for i in range(n)])
    with nb.Context(0) as c:
        for kw in (dict(blur=1, blur_radius=21, edge_detect=1),                            # fused kernel per chunk
                   dict(invert=1, blur=1, blur_radius=7, sharpen=1, edge_detect=1)):       # four unfused stages per chunk
            f = uniform(w, h, **kw)
            eo, es = oracle.chain_batch(frames, oracle_uniform(oracle, f), oracle.CS_RAW)
            out = np.empty_like(frames)
            s = c.chain_host(f, frames, out, n, nb.NORM_CS_RAW)
            assert np.array_equal(out, eo) and np.abs(s - es).max() <= SSIM_TIGHT, kw
            src, dst = c.image_from_host(frames), c.image_create(w, h, n)
            s = c.chain(f, src, dst, nb.NORM_CS_RAW)
            assert np.array_equal(dst.download(), eo) and np.abs(s - es).max() <= SSIM_TIGHT, kw
            src.destroy(), dst.destroy()


def test_uniform_validation(ctx):
    src, dst = ctx.image_from_host(splitmix_rgba(16, 8)), ctx.image_create(16, 8, 1)
    with pytest.raises(nb.NormError) as e:
        ctx.chain(uniform(8, 8), src, dst)
    assert e.value.status == nb._lib.NORM_E_DIM_MISMATCH
    with pytest.raises(nb.NormError) as e:
        ctx.chain(uniform(16, 8, blur=1, blur_radius=0), src, dst)
    assert e.value.status == nb._lib.NORM_E_INVALID_ARG
    bad = ctx.image_create(16, 9, 1)
    with pytest.raises(nb.NormError):
        ctx.effects(uniform(16, 8), src, bad)
    src.destroy(), dst.destroy(), bad.destroy()


# ------------------------------------------------------------------ BASELINE configs

def test_config1_obama_luma_blur(ctx, oracle):
    """configs[0]: tests/obama.png -> compat expansion (alpha 0) -> grayscale + blur r=21, sRGB texture."""
    d = read_png(oracle, os.path.join(GOLDEN, "images", "obama.png"))
    png = to_png(d)
    tex = nb.Texture.from_image(ctx, png, nb.NORM_EXPAND_COMPAT)
    rgba = tex.image.download()[0]
    exp_rgba = oracle.expand_rgba8(d, oracle.EXPAND_COMPAT)
    assert np.array_equal(rgba, exp_rgba) and (rgba[..., 3] == 0).all()
    f = nb.FeatureUniform(png.width, png.height, png.gamma())
    f.toggle_grayscale(), f.toggle_blur()
    pipe = nb.reference_pipeline(ctx, tex, f)
    out = pipe.render(f)
    got = out.download()[0]
    exp = oracle.effects(exp_rgba, oracle_uniform(oracle, f), oracle.CS_REFERENCE)
    assert np.array_equal(got, exp)
    # the staged EffectPipeline::execute leaves the same texture A the fused front stage computes
    pipe.execute(f)
    assert np.array_equal(pipe.texture_a().image.download()[0], oracle.front_s(exp_rgba, oracle.CS_REFERENCE, 0, 1, 0))
    # literal shader loop order on a crop: within 1 LSB
    crop = oracle.front_s(exp_rgba, oracle.CS_REFERENCE, 0, 1, 0)[200:328, 300:428].copy()
    gc = run_stage(ctx, lambda s, d_: ctx.blur(21, s, d_), crop)[0]
    assert np.abs(gc.astype(int) - oracle.blur_r(crop, 21).astype(int)).max() <= 1
    # SSIM of the effect against the source, through the Png-level API
    s = ctx.ssim(out, tex.image)[0]
    assert abs(s - oracle.ssim(exp, exp_rgba)) <= SSIM_TIGHT
    out.destroy(), tex.destroy()


def test_config2_pngsuite_sobel(ctx, oracle):
    """configs[1]: every valid PngSuite file: spec expansion -> Sobel line detection."""
    for name in VALID:
        d = read_png(oracle, os.path.join(PNGSUITE, name + ".png"))
        tex = nb.Texture.from_image(ctx, to_png(d), nb.NORM_EXPAND_SPEC)
        f = nb.FeatureUniform(d.width, d.height, d.gamma)
        f.toggle_edge_detect()
        out = ctx.image_create(d.width, d.height, 1)
        ctx.effects(f, tex.image, out)
        exp = oracle.effects(oracle.expand_rgba8(d, oracle.EXPAND_SPEC), oracle_uniform(oracle, f), oracle.CS_REFERENCE)
        assert np.array_equal(out.download()[0], exp), name
        out.destroy(), tex.destroy()


def periodic_table_8k(oracle):
    """BASELINE.md config 3, real content: the reference's tests/Periodic_table_large.png (6000x3300 RGBA8, decoded by
    the oracle's PNG front-end) tiled with wrap-around to 7680x4320."""
    d = oracle.png_decode(open(os.path.join(GOLDEN, "images", "Periodic_table_large.png"), "rb").read())
    tile = oracle.expand_rgba8(d, oracle.EXPAND_SPEC)
    assert tile.shape == (3300, 6000, 4)
    ys, xs = np.arange(4320) % 3300, np.arange(7680) % 6000
    return np.ascontiguousarray(tile[ys][:, xs])


@pytest.mark.parametrize("content", ["prng", "periodic_table"])
def test_config3_8k_chain_properties(ctx, oracle, content):
    """configs[2] at full size (7680x4320), on the PRNG image and on the tiled periodic table, for every radius of the
    BASELINE sweep: oracle on sampled row bands (top edge, middle, bottom edge) + size-independent properties."""
    w, h = 7680, 4320
    if content == "prng":
        img = splitmix_rgba(w, h, alpha=255)
        yy = (np.arange(h, dtype=np.int64)[:, None] * 255 // (h - 1)).astype(np.uint8)
        img[..., 0] = (img[..., 0].astype(np.uint16) + yy) // 2   # gradient so blur output is not flat
    else:
        img = periodic_table_8k(oracle)
    src, dst = ctx.image_from_host(img), ctx.image_create(w, h, 1)
    for r in (3, 7, 21, 39):
        f = uniform(w, h, grayscale=1, blur=1, blur_radius=r, edge_detect=1)
        s = ctx.chain(f, src, dst, nb.NORM_CS_RAW)[0]
        got = dst.download()[0]
        # (1) bands incl. top and bottom edges vs the oracle run on the band + its stencil halo
        halo = (r - 1) // 2 + 2
        for y0, y1 in ((0, 64), (2100, 2164), (h - 64, h)):
            lo, hi = max(0, y0 - halo), min(h, y1 + halo)
            band = np.ascontiguousarray(img[lo:hi])
            eo, _ = oracle.chain(band, oracle.make_uniform(w, hi - lo, grayscale=1, blur=1, blur_radius=r, edge_detect=1), oracle.CS_RAW)
            assert np.array_equal(got[y0:y1], eo[y0 - lo:y1 - lo]), (content, r, y0, y1)
        # (2) output structure: gray, opaque
        assert (got[..., 3] == 255).all() and np.array_equal(got[..., 0], got[..., 1]) and np.array_equal(got[..., 0], got[..., 2])
        # (3) SSIM of the chain output vs source equals the oracle's SSIM on the GPU output
        assert abs(s - oracle.ssim(got, img)) <= SSIM_TIGHT, (content, r)
    # (4) SSIM(x,x) == 1
    assert abs(ctx.ssim(src, src)[0] - 1.0) <= 1e-12
    src.destroy(), dst.destroy()


def test_config4_batch_frames_sample(ctx, oracle):
    """configs[3] geometry (1920x1080 frames, blur r=21 + Sobel + SSIM) on a 6-frame sample;
    frame results must not depend on batch position."""
    w, h = 1920, 1080
    frames = np.stack([splitmix_rgba(w, h, seed=0x6E6F726D ^ (i << 20), alpha=255) // 2 + smooth_rgba(w, h, seed=i) // 2
                       for i in range(3)])
    frames = np.concatenate([frames, frames[::-1]])
    f = uniform(w, h, blur=1, blur_radius=21, edge_detect=1)
    src, dst = ctx.image_from_host(frames), ctx.image_create(w, h, 6)
    s = ctx.chain(f, src, dst, nb.NORM_CS_RAW)
    got = dst.download()
    eo, es = oracle.chain_batch(frames[:3], oracle_uniform(oracle, f), oracle.CS_RAW)
    assert np.array_equal(got[:3], eo) and np.abs(s[:3] - es).max() <= SSIM_TIGHT
    assert np.array_equal(got[3:], got[:3][::-1]) and np.array_equal(s[3:], s[:3][::-1])
    src.destroy(), dst.destroy()


def test_fused_sqrt_exhaustive_intensity_radicands(ctx, oracle):
    """The fused kernel replaces sqrtf by the in-range Newton sequence: every possible intensity
    radicand r^2+g^2+b^2 must still round exactly as sqrtf does.  4x4 constant blocks survive an
    r=3 blur unchanged in their 2x2 interior, so every distinct radicand reaches the I pass."""
    v = np.arange(256, dtype=np.int64)
    r, g, b = np.meshgrid(v, v, v, indexing="ij")
    keep = (r <= g) & (g <= b)
    n = (r * r + g * g + b * b)[keep]
    _, first = np.unique(n, return_index=True)
    cols = np.stack([r[keep][first], g[keep][first], b[keep][first]], -1).astype(np.uint8)
    nb_cols = cols.shape[0]
    side = int(np.ceil(np.sqrt(nb_cols)))
    grid = np.zeros((side * side, 4), np.uint8)
    grid[:nb_cols, :3] = cols
    grid[:, 3] = 255
    img = np.repeat(np.repeat(grid.reshape(side, side, 4), 4, axis=0), 4, axis=1).copy()
    h, w, _ = img.shape
    f = uniform(w, h, blur=1, blur_radius=3, edge_detect=1)
    src, dst = ctx.image_from_host(img), ctx.image_create(w, h, 1)
    s = ctx.chain(f, src, dst, nb.NORM_CS_RAW)[0]
    eo, es = oracle.chain(img, oracle_uniform(oracle, f), oracle.CS_RAW)
    assert np.array_equal(dst.download()[0], eo) and abs(s - es) <= SSIM_TIGHT
    src.destroy(), dst.destroy()


def test_fused_matches_unfused_large_random(ctx, oracle):
    """Dense random Sobel inputs: fused kernel vs the stage kernels (which use sqrtf) vs oracle."""
    w, h = 1531, 1187
    img = splitmix_rgba(w, h, seed=99)
    f = uniform(w, h, blur=1, blur_radius=7, edge_detect=1)
    src, dst = ctx.image_from_host(img), ctx.image_create(w, h, 1)
    ctx.chain(f, src, dst, nb.NORM_CS_RAW)
    fused = dst.download()[0]
    tmp = ctx.image_create(w, h, 1)
    ctx.blur(7, src, tmp)
    ctx.sobel(tmp, dst)
    assert np.array_equal(fused, dst.download()[0])
    assert np.array_equal(fused, oracle.sobel_s(oracle.blur_s(img, 7)))
    src.destroy(), dst.destroy(), tmp.destroy()


@pytest.mark.parametrize("nranks", [2, 3, 5])
def test_band_mode_emulated_ranks(ctx, oracle, nranks):
    """configs[4] logic on one GPU: each emulated rank runs norm_cuda_chain_band on its band + halo
    (halo rows arrive through norm_cuda_copy_rows from the 'peer' band buffers); owned rows and the
    summed SSIM moments must equal the single-image chain and the oracle."""
    w, h = 300, 211
    img = smooth_rgba(w, h, seed=31)
    for kw in (dict(blur=1, blur_radius=21, edge_detect=1), dict(grayscale=1, blur=1, blur_radius=7, edge_detect=1),
               dict(blur=1, blur_radius=5, sharpen=1), dict(blur=1, blur_radius=3, edge_detect=1)):
        f = uniform(w, h, **kw)
        lays = [nb.sharding.band_layout(f, h, nranks, r) for r in range(nranks)]
        bands = [ctx.image_create(w, l["ext_rows"], 1) for l in lays]
        outs = [ctx.image_create(w, l["ext_rows"], 1) for l in lays]
        for l, bimg in zip(lays, bands):  # each rank starts with its OWN rows only (halo rows poisoned)
            ext = np.full((l["ext_rows"], w, 4), 0xAB, np.uint8)
            ext[l["own_off"]:l["own_off"] + l["rows"]] = img[l["row0"]:l["row0"] + l["rows"]]
            bimg.upload(ext)
        for r, bimg in enumerate(bands):  # halo exchange
            for (peer, prow, mine, n) in nb.sharding.halo_requests(f, h, nranks, r):
                ctx.copy_rows(bimg, mine, bands[peer].device_ptr, bands[peer].pitch, prow, n)
        ctx.sync()
        total = np.zeros(5)
        rows = []
        for l, bimg, oimg in zip(lays, bands, outs):
            fb = uniform(w, l["ext_rows"], **kw)
            total += ctx.chain_band(fb, bimg, oimg, l["own_off"], l["rows"], nb.NORM_CS_RAW)[0]
            rows.append(oimg.download()[0][l["own_off"]:l["own_off"] + l["rows"]])
        got = np.concatenate(rows)
        eo, es = oracle.chain(img, oracle_uniform(oracle, f), oracle.CS_RAW)
        assert np.array_equal(got, eo), (nranks, kw)
        assert abs(nb.ssim_from_moments(total, w * h) - es) <= SSIM_TIGHT
        for x in bands + outs:
            x.destroy()


@pytest.mark.parametrize("nranks", [1, 2, 3, 5])
def test_band_scheduler_emulated_ranks(oracle, nranks):
    """configs[4] through norm_cuda_band_*: one band object per emulated rank (one context each, all on this
    GPU, handles exchanged inside the process), halo rows by peer copies on the halo stream, interior rows
    launched before the halo rows arrive, edge strips after; owned rows and the summed SSIM moments must equal
    the oracle on the whole image.  Two steps in a row with changed source rows exercise begin_update / publish."""
    w = 300
    for kw, h in ((dict(blur=1, blur_radius=21, edge_detect=1), 211), (dict(grayscale=1, blur=1, blur_radius=7, edge_detect=1), 97),
                  (dict(blur=1, blur_radius=5, sharpen=1), 64), (dict(blur=1, blur_radius=21, edge_detect=1), 40),
                  (dict(blur=1, blur_radius=3, edge_detect=1), 131)):
        f = uniform(w, h, **kw)
        ctxs = [nb.Context(0) for _ in range(nranks)]
        bands = [nb.Band(c, f, w, h, nranks, r) for r, c in enumerate(ctxs)]
        handles = [b.export() for b in bands]
        for b in bands:
            b.connect(handles)
        moms = [c.image_create(64, 1, 1) for c in ctxs]  # 256 bytes of device memory each
        for step, seed in enumerate((31, 32)):
            img = smooth_rgba(w, h, seed=seed)
            for b in bands:
                b.begin_update()
                ext = np.full((b.src.height, w, 4), 0xAB, np.uint8)  # halo rows poisoned: they must come from the peers
                ext[b.halo_up:b.halo_up + b.rows] = img[b.row0:b.row0 + b.rows]
                b.src.upload(ext)
                b.publish()
            for b, m in zip(bands, moms):  # (all ranks have published: the in-process stand-in for the host barrier)
                b.step_async(nb.NORM_CS_RAW, m.device_ptr)
            for b in bands:
                b.sync()
            got = np.concatenate([b.own_rows_of(b.dst) for b in bands])
            total = sum(np.frombuffer(m.download()[0].tobytes()[:40], np.float64) for m in moms)
            eo, es = oracle.chain(img, oracle_uniform(oracle, f), oracle.CS_RAW)
            assert np.array_equal(got, eo), (nranks, kw, h, step)
            assert abs(nb.ssim_from_moments(total, w * h) - es) <= SSIM_TIGHT, (nranks, kw, h, step)
        for m in moms:
            m.destroy()
        for b in bands:
            b.destroy()
        for c in ctxs:
            c.close()


def test_effects_flips(ctx, oracle):
    """FlipX / FlipY (feature_uniform.rs:143-175): the finished frame is mirrored."""
    for (w, h) in [(5, 7), (130, 70), (333, 41)]:
        img = smooth_rgba(w, h, seed=w)
        for fx, fy in ((1, 0), (0, 1), (1, 1)):
            for kw in (dict(), dict(blur=1, blur_radius=7), dict(grayscale=1, edge_detect=1)):
                f = uniform(w, h, **kw)
                if fx:
                    f.apply_transform("FlipX")
                if fy:
                    f.apply_transform("FlipY")
                got = run_stage(ctx, lambda s, d: ctx.effects(f, s, d, nb.NORM_CS_RAW), img)[0]
                f0 = uniform(w, h, **kw)
                base = oracle.effects(img, oracle_uniform(oracle, f0), oracle.CS_RAW)
                exp = base[::-1 if fy else 1, ::-1 if fx else 1]
                assert np.array_equal(got, exp), (w, h, fx, fy, kw)
                of = oracle_uniform(oracle, f)
                of.transform[0][0], of.transform[1][1] = f.raw.transform[0][0], f.raw.transform[1][1]
                assert np.array_equal(oracle.effects(img, of, oracle.CS_RAW), exp)


# every radius the reference's UI can reach: 3..39 in steps of 2 (feature_uniform.rs:66-84)
@pytest.mark.parametrize("radius", list(range(3, 41, 2)))
def test_every_ui_radius_chain_and_blur(ctx, oracle, radius):
    w, h = 251, 97
    img = smooth_rgba(w, h, seed=radius) if radius % 4 == 1 else splitmix_rgba(w, h, seed=radius)
    for kw in (dict(blur=1, blur_radius=radius, edge_detect=1), dict(grayscale=1, blur=1, blur_radius=radius, edge_detect=1),
               dict(gamma=45455, blur=1, blur_radius=radius, edge_detect=1)):
        f = uniform(w, h, **kw)
        src, dst = ctx.image_from_host(img), ctx.image_create(w, h, 1)
        ssim = ctx.chain(f, src, dst, nb.NORM_CS_REFERENCE)
        exp, es = oracle.chain_batch(img[None], oracle_uniform(oracle, f), nb.NORM_CS_REFERENCE)
        assert np.array_equal(dst.download(), exp), (radius, kw)
        assert abs(ssim[0] - es[0]) <= SSIM_TIGHT, (radius, kw)
        src.destroy(), dst.destroy()
    for kw in (dict(blur=1, blur_radius=radius), dict(grayscale=1, blur=1, blur_radius=radius)):
        f = uniform(w, h, **kw)
        got = run_stage(ctx, lambda s, d: ctx.effects(f, s, d, nb.NORM_CS_RAW), img)[0]
        assert np.array_equal(got, oracle.effects(img, oracle_uniform(oracle, f), nb.NORM_CS_RAW)), (radius, kw)
