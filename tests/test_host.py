"""CPU-only checks of the boundary: the C-ABI library loads and exports every declared symbol,
host-side planning helpers agree with the oracle, and the mirrored FeatureUniform behaves
like src/renderer/feature_uniform.rs.  No compute entry point is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import norm_b200 as nb
from conftest import ROOT


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "norm_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(norm_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"norm_status"}
    assert len(declared) >= 35
    lib = C.CDLL(nb.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in norm_cuda.h but not exported"
    assert declared == set(nb.PROTOTYPES), declared ^ set(nb.PROTOTYPES)
    assert nb.load().norm_cuda_abi_version() == 1


def test_rust_sys_crate_declares_every_exported_symbol():
    """rust/norm-cuda-sys/src/ffi.rs is generated from the headers (tools/gen_rust_ffi.py): it must be current, and
    the symbols it declares must be exactly the norm_* functions the two libraries export (nm -D)."""
    import subprocess
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_rust_ffi
    text, names = gen_rust_ffi.render()
    assert open(gen_rust_ffi.OUT).read() == text, "run `python tools/gen_rust_ffi.py`"
    exported = set()
    for lib in (nb.LIB_PATH, os.path.join(os.path.dirname(nb.LIB_PATH), "libnorm_nccl.so")):
        out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True, check=True).stdout
        exported |= {ln.split()[-1] for ln in out.splitlines() if ln.split()[-2:-1] == ["T"] and ln.split()[-1].startswith("norm_")}
    assert set(names) == exported, set(names) ^ exported
    # the hand-written structs carry the sizes the C side asserts
    rs = open(os.path.join(ROOT, "rust", "norm-cuda-sys", "src", "lib.rs")).read()
    for name, size in (("NormFeatureUniform", 112), ("NormPngDesc", C.sizeof(nb._lib.NormPngDesc)), ("NormBandHandle", 256),
                       ("NormInflatedDesc", C.sizeof(nb._lib.NormInflatedDesc)), ("NormDrawUniform", 80)):
        assert re.search(r"size_of::<%s>\(\) == %d" % (name, size), rs), (name, size)


def test_no_oracle_or_cpu_fallback_in_product():
    """The product must not reach into oracle/ (task rule 3)."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "norm_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert not re.search(r"#\s*include[^\n]*oracle|import[^\n]*oracle|libnorm_oracle|oracle_lib|dlopen", src), fn


def test_feature_uniform_layout_and_methods():
    assert C.sizeof(nb.NormFeatureUniform) == 112
    offs = {n: getattr(nb.NormFeatureUniform, n).offset for n, _ in nb.NormFeatureUniform._fields_}
    assert [offs[k] for k in ("grayscale", "invert", "gamma", "blur", "blur_radius", "width", "height", "sharpen",
                              "sharpen_factor", "edge_detect", "_padding_1", "transform")] == \
        [0, 4, 8, 12, 16, 20, 24, 28, 32, 36, 40, 48]
    f = nb.FeatureUniform(640, 480, 45455)
    assert (f.raw.blur_radius, f.raw.sharpen_factor, f.gamma()) == (21, 16, 45455)
    assert not f.blur()
    f.toggle_blur()
    assert f.blur()
    for _ in range(20):
        f.increase_blur_radius()
    assert f.raw.blur_radius == 39
    for _ in range(40):
        f.decrease_blur_radius()
    assert f.raw.blur_radius == 3
    for _ in range(50):
        f.increase_sharpen_factor()
    assert f.raw.sharpen_factor == 40
    f.apply_transform("FlipX")
    assert f.raw.transform[0][0] == -1.0 and f.raw.transform[1][1] == 1.0
    f.apply_transform("FlipX")
    assert f.raw.transform[0][0] == 1.0


def test_host_tables_match_oracle(oracle):
    for r in (1, 2, 3, 4, 7, 21, 39, 63):
        t, d = nb.blur_taps(r)
        ot, od = oracle.blur_taps(r)
        assert d == od and np.array_equal(t, ot)
    with pytest.raises(nb.NormError):
        nb.blur_taps(0)
    with pytest.raises(nb.NormError):
        nb.blur_taps(64)
    for cs in (0, 1):
        for g in (0, 35000, 45455, 100000, 250000):
            assert np.array_equal(nb.front_lut(cs, g), oracle.front_lut(cs, g))
    m = np.array([1000.0, 1100.0, 120000.0, 150000.0, 130000.0])
    assert nb.ssim_from_moments(m, 10) == oracle.ssim_from_moments(m, 10)


def test_plans():
    # frames: contiguous, complete, balanced
    for frames, n in ((4096, 8), (10, 3), (2, 4)):
        got = [nb.plan_frames(frames, n, r) for r in range(n)]
        assert got[0][0] == 0 and sum(c for _, c in got) == frames
        assert all(got[i][0] + got[i][1] == got[i + 1][0] for i in range(n - 1))
        assert max(c for _, c in got) - min(c for _, c in got) <= 1
    f = nb.FeatureUniform(32768, 32768)
    f.toggle_blur()
    f.toggle_edge_detect()
    plans = [nb.plan_band(f, 32768, 8, r) for r in range(8)]
    assert plans[0] == (0, 4096, 0, 11) and plans[3] == (12288, 4096, 11, 11) and plans[7] == (28672, 4096, 11, 0)
    f.raw.blur_radius = 4  # even radius: asymmetric footprint
    assert nb.plan_band(f, 100, 2, 0) == (0, 50, 0, 2) and nb.plan_band(f, 100, 2, 1) == (50, 50, 3, 0)


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(nb.NormError) as e:
        nb.Context(0)
    assert e.value.status == nb._lib.NORM_E_CUDA and "no CPU fallback" in str(e.value)


def test_shape_structs_and_srgb_tables(oracle):
    """shape_uniform.rs / draw_uniform.rs layouts and the host-side sRGB tables of the overlay."""
    import ctypes as C
    from norm_b200 import _lib, api
    assert C.sizeof(_lib.NormShapeUniform) == 16 and C.sizeof(_lib.NormCircleData) == 16
    assert C.sizeof(_lib.NormDrawUniform) == 80 and _lib.NormDrawUniform.camera_view_proj.offset == 16
    su = api.ShapeUniform(640, 480)
    assert su.raw.selected_circle == 0xFFFFFFFF and su.raw.num_circles == 0
    su.set_selected_circle(3), su.set_num_circles(7), su.update_dimensions(10, 20)
    assert (su.raw.width, su.raw.height, su.raw.num_circles, su.raw.selected_circle) == (10, 20, 7, 3)
    su.set_selected_circle(None)
    assert su.raw.selected_circle == 0xFFFFFFFF
    # Circle round trip pixel <-> normalised space (shape.rs:188-243)
    c = api.Circle.from_pixel_space((320.0, 120.0), 48.0, (640.0, 480.0))
    assert c.center == (0.5, 0.25) and c.radius == np.float32(0.1)
    p = c.in_pixel_space((640.0, 480.0))
    assert p.center == (320.0, 120.0) and abs(p.radius - 48.0) < 1e-4
    d = c.data()
    assert (d.x, d.y, d._padding) == (0.5, 0.25, 0.0)
    du = api.DrawUniform()
    assert not du.crosshair() and du.raw.camera_view_proj[2][2] == 1.0
    du.toggle_crosshair()
    assert du.crosshair()
    e, t = api.srgb_tables()
    oe, ot = oracle.srgb_tables()
    assert (e == oe).all() and (t == ot).all()
