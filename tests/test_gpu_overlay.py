"""Box blur (image_shader.wgsl:105-118) and the shape overlay / composite / crosshair
(shape_shader.wgsl:41-89, image_shader.wgsl:176-193) on the GPU, bit-exact against the
CPU oracle."""
import numpy as np
import pytest

import norm_b200 as nb
from norm_b200 import api
from conftest import smooth_rgba, splitmix_rgba

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = nb.Context(0)
    yield c
    c.close()


def run_stage(ctx, img, fn):
    src = ctx.image_from_host(img)
    dst = ctx.image_create(src.width, src.height, src.frames)
    try:
        fn(src, dst)
        return dst.download().reshape(img.shape)
    finally:
        src.destroy(), dst.destroy()


@pytest.mark.parametrize("w,h", [(1, 1), (3, 2), (127, 17), (128, 16), (129, 33), (641, 359), (1920, 1080)])
def test_box_blur_bit_exact(ctx, oracle, w, h):
    img = splitmix_rgba(w, h, seed=w * 7 + h)
    got = run_stage(ctx, img, ctx.box_blur)
    assert (got == oracle.box_blur_s(img)).all()


def random_circles(rng, n, big=False):
    out = []
    for _ in range(n):
        r = rng.uniform(0.002, 0.3 if big else 0.08)
        out.append((float(np.float32(rng.uniform(-0.1, 1.1))), float(np.float32(rng.uniform(-0.1, 1.1))),
                    float(np.float32(r))))
    return out


def gpu_overlay(ctx, img, circles, selected, crosshair):
    draw = None
    if crosshair is not None:
        draw = api.DrawUniform()
        draw.toggle_crosshair()
        draw.set_circle_center(crosshair[0], crosshair[1])
        draw.set_circle_radius(crosshair[2])
    cs = [api.Circle((x, y), r) for x, y, r in circles]
    return run_stage(ctx, img, lambda s, d: ctx.overlay(s, d, cs, selected, draw))


@pytest.mark.parametrize("w,h,n,selected,cross", [
    (64, 48, 0, None, None),
    (64, 48, 1, 0, None),
    (160, 120, 4, 1, (20.0, 100.0, 6.5)),
    (333, 211, 40, 17, (300.5, 10.25, 33.0)),
    (129, 9, 7, None, (0.0, 0.0, 4.0)),
    (5, 3, 3, 2, (2.5, 1.5, 1.0)),
    (1920, 1080, 256, 255, (960.0, 540.0, 25.0)),
])
def test_overlay_bit_exact(ctx, oracle, w, h, n, selected, cross):
    rng = np.random.default_rng(w * 31 + n)
    img = splitmix_rgba(w, h, seed=n + 3)
    circles = random_circles(rng, n, big=n <= 8)
    got = gpu_overlay(ctx, img, circles, selected, cross)
    exp = oracle.overlay(img, circles, selected, cross)
    assert (got == exp).all(), f"{(got != exp).any(axis=-1).sum()} pixels differ"
    if n:
        assert (exp != img).any()  # the case exercises something


def test_overlay_degenerate_parameters(ctx, oracle):
    img = smooth_rgba(96, 64, seed=4)
    circles = [(0.5, 0.5, 0.0), (0.25, 0.25, -0.1), (float("nan"), 0.5, 0.1), (0.5, 0.5, float("inf")),
               (0.75, 0.75, 1e-4), (0.1, 0.9, 3.0e7)]
    for sel in (None, 0, 1, 2, 3, 5):
        got = gpu_overlay(ctx, img, circles[:3] + circles[4:5], sel if sel in (None, 0, 1, 2) else None, None)
        exp = oracle.overlay(img, circles[:3] + circles[4:5], sel if sel in (None, 0, 1, 2) else None, None)
        assert (got == exp).all()
    got = gpu_overlay(ctx, img, circles, 3, (48.0, 32.0, 1e9))
    exp = oracle.overlay(img, circles, 3, (48.0, 32.0, 1e9))
    assert (got == exp).all()


def test_overlay_multi_frame_and_order(ctx, oracle):
    frames = np.stack([splitmix_rgba(200, 90, seed=s) for s in (1, 2, 3)])
    circles = [(0.5, 0.5, 0.3), (0.52, 0.5, 0.25), (0.2, 0.3, 0.1)]
    got = gpu_overlay(ctx, frames, circles, 0, (100.0, 45.0, 10.0))
    for i in range(3):
        assert (got[i] == oracle.overlay(frames[i], circles, 0, (100.0, 45.0, 10.0))).all()
    # list order matters for the outline: selecting a circle drawn under a later one
    a = oracle.overlay(frames[0], circles, 0)
    b = oracle.overlay(frames[0], [circles[1], circles[0], circles[2]], 1)
    assert (a != b).any()


def test_overlay_errors(ctx):
    img = splitmix_rgba(32, 32, seed=1)
    src = ctx.image_from_host(img)
    dst = ctx.image_create(32, 32, 1)
    with pytest.raises(nb.NormError):
        ctx.overlay(src, dst, [api.Circle((0.5, 0.5), 0.1)] * 257)
    small = ctx.image_create(16, 32, 1)
    with pytest.raises(nb.NormError):
        ctx.overlay(src, small, [])
    with pytest.raises(nb.NormError):
        ctx.overlay(src, src, [])
    src.destroy(), dst.destroy(), small.destroy()
