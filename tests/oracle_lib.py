"""ctypes view of oracle/libnorm_oracle.so -- the CPU restatement of norm's hot path.

Test infrastructure: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs only.  Never imported by the norm_b200 package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ORACLE_DIR = os.path.join(_ROOT, "oracle")
_SO = os.path.join(_ORACLE_DIR, "libnorm_oracle.so")

OK, E_INVALID, E_UNSUPPORTED, E_CORRUPT = 0, 1, 2, 3
EXPAND_COMPAT, EXPAND_SPEC = 0, 1
CS_REFERENCE, CS_RAW = 0, 1


class OraclePng(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("bit_depth", C.c_uint8), ("color_type", C.c_uint8), ("compression", C.c_uint8),
        ("filter", C.c_uint8), ("interlace", C.c_uint8),
        ("gamma", C.c_uint32),
        ("pixels", C.POINTER(C.c_uint8)), ("len", C.c_size_t),
        ("plte", C.c_uint8 * 768), ("n_plte", C.c_int32),
        ("trns", C.c_uint8 * 256), ("n_trns", C.c_int32),
    ]


class FeatureUniform(C.Structure):
    """feature_uniform.rs:1-17, 112 bytes."""
    _fields_ = [
        ("grayscale", C.c_uint32), ("invert", C.c_uint32), ("gamma", C.c_uint32),
        ("blur", C.c_uint32), ("blur_radius", C.c_uint32), ("width", C.c_uint32),
        ("height", C.c_uint32), ("sharpen", C.c_uint32), ("sharpen_factor", C.c_uint32),
        ("edge_detect", C.c_uint32), ("pad", C.c_uint8 * 8), ("transform", (C.c_float * 4) * 4),
    ]


assert C.sizeof(FeatureUniform) == 112


class ShapeUniform(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("num_circles", C.c_uint32),
                ("selected_circle", C.c_uint32)]


class CircleData(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("radius", C.c_float), ("pad", C.c_float)]


class DrawUniform(C.Structure):
    _fields_ = [("crosshair", C.c_uint32), ("circle_center_x", C.c_float), ("circle_center_y", C.c_float),
                ("circle_radius", C.c_float), ("camera_view_proj", (C.c_float * 4) * 4)]


def make_uniform(w, h, gamma=0, grayscale=0, invert=0, blur=0, blur_radius=21, sharpen=0,
                 sharpen_factor=16, edge_detect=0):
    f = FeatureUniform()
    f.grayscale, f.invert, f.gamma, f.blur, f.blur_radius = grayscale, invert, gamma, blur, blur_radius
    f.width, f.height, f.sharpen, f.sharpen_factor, f.edge_detect = w, h, sharpen, sharpen_factor, edge_detect
    for i in range(4):
        f.transform[i][i] = 1.0
    return f


def build(force=False):
    src = [os.path.join(_ORACLE_DIR, n) for n in ("norm_oracle.c", "norm_oracle.h", "Makefile")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-C", _ORACLE_DIR, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        u8p, f32p, f64p = C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_double)
        L.oracle_png_decode.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(OraclePng)]
        L.oracle_png_free.argtypes = [C.POINTER(OraclePng)]
        L.oracle_adam7_pass_sizes.argtypes = [C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.oracle_row_bytes.argtypes = [C.c_uint32, C.c_int, C.c_int]
        L.oracle_row_bytes.restype = C.c_size_t
        L.oracle_filter_scanlines.argtypes = [u8p, C.c_uint32, C.c_size_t, C.c_size_t, C.c_int, u8p, u8p]
        L.oracle_filter_scanlines.restype = None
        L.oracle_expand_rgba8.argtypes = [C.POINTER(OraclePng), C.c_int, u8p]
        L.oracle_rgb8.argtypes = [C.POINTER(OraclePng), u8p]
        L.oracle_front_lut.argtypes = [C.c_int, C.c_uint32, u8p]
        L.oracle_front_lut.restype = None
        L.oracle_front_s.argtypes = [u8p, C.c_uint32, C.c_uint32, u8p, C.c_int, C.c_int, u8p]
        L.oracle_front_s.restype = None
        L.oracle_front_r.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, C.c_int, C.c_int, u8p]
        L.oracle_front_r.restype = None
        L.oracle_blur_taps.argtypes = [C.c_uint32, f32p, C.POINTER(C.c_int)]
        for n in ("oracle_blur_s", "oracle_blur_r"):
            getattr(L, n).argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p]
        for n in ("oracle_sobel_s", "oracle_sobel_r"):
            getattr(L, n).argtypes = [u8p, C.c_uint32, C.c_uint32, u8p]
            getattr(L, n).restype = None
        for n in ("oracle_box_blur_s", "oracle_box_blur_r"):
            getattr(L, n).argtypes = [u8p, C.c_uint32, C.c_uint32, u8p]
            getattr(L, n).restype = None
        L.oracle_srgb_tables.argtypes = [f32p, f32p]
        L.oracle_srgb_tables.restype = None
        L.oracle_srgb_encode_lut.argtypes = [u8p]
        L.oracle_srgb_encode_lut.restype = None
        L.oracle_encode_srgb.argtypes = [u8p, C.c_size_t, u8p]
        L.oracle_encode_srgb.restype = None
        L.oracle_overlay.argtypes = [u8p, C.c_uint32, C.c_uint32, C.POINTER(ShapeUniform), C.POINTER(CircleData),
                                     C.POINTER(DrawUniform), u8p]
        for n in ("oracle_sharpen_s", "oracle_sharpen_r"):
            getattr(L, n).argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p]
            getattr(L, n).restype = None
        L.oracle_ssim_moments.argtypes = [u8p, u8p, C.c_size_t, f64p]
        L.oracle_ssim_moments.restype = None
        L.oracle_ssim_from_moments.argtypes = [f64p, C.c_double]
        L.oracle_ssim_from_moments.restype = C.c_double
        L.oracle_ssim.argtypes = [u8p, u8p, C.c_size_t]
        L.oracle_ssim.restype = C.c_double
        L.oracle_ssim_luma.argtypes = [u8p, u8p, C.c_size_t, C.c_int, C.c_int]
        L.oracle_ssim_luma.restype = C.c_double
        L.oracle_ssim_f32_literal.argtypes = [u8p, u8p, C.c_size_t]
        L.oracle_ssim_f32_literal.restype = C.c_float
        L.oracle_effects.argtypes = [C.POINTER(FeatureUniform), C.c_int, u8p, u8p]
        L.oracle_chain.argtypes = [C.POINTER(FeatureUniform), C.c_int, u8p, u8p, f64p]
        L.oracle_chain_batch.argtypes = [C.POINTER(FeatureUniform), C.c_int, u8p, u8p, f64p, C.c_int]
        L.oracle_num_threads.restype = C.c_int
        L.oracle_set_num_threads.argtypes = [C.c_int]
        L.oracle_set_num_threads.restype = None
        _lib = L
    return _lib


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


class DecodedPng:
    """Plain-Python copy of an oracle_png (pixels as a numpy array)."""

    def __init__(self, p: OraclePng):
        self.width, self.height = p.width, p.height
        self.bit_depth, self.color_type = p.bit_depth, p.color_type
        self.compression, self.filter, self.interlace = p.compression, p.filter, p.interlace
        self.gamma = p.gamma
        self.pixels = np.ctypeslib.as_array(p.pixels, shape=(p.len,)).copy() if p.len else np.zeros(0, np.uint8)
        self.plte = np.frombuffer(bytes(p.plte), np.uint8)[: 3 * p.n_plte].copy()
        self.trns = np.frombuffer(bytes(p.trns), np.uint8)[: p.n_trns].copy()

    def as_struct(self):
        p = OraclePng()
        p.width, p.height = self.width, self.height
        p.bit_depth, p.color_type = self.bit_depth, self.color_type
        p.compression, p.filter, p.interlace, p.gamma = self.compression, self.filter, self.interlace, self.gamma
        self._keep = np.ascontiguousarray(self.pixels)
        p.pixels = _u8(self._keep)
        p.len = self._keep.size
        C.memmove(p.plte, self.plte.tobytes(), self.plte.size)
        p.n_plte = self.plte.size // 3
        C.memmove(p.trns, self.trns.tobytes(), self.trns.size)
        p.n_trns = self.trns.size
        return p


def png_decode(data: bytes):
    """-> DecodedPng, raises ValueError for files the front-end rejects."""
    p = OraclePng()
    rc = lib().oracle_png_decode(data, len(data), C.byref(p))
    if rc != OK:
        raise ValueError(f"oracle_png_decode rc={rc}")
    try:
        return DecodedPng(p)
    finally:
        lib().oracle_png_free(C.byref(p))


def expand_rgba8(png: DecodedPng, mode):
    out = np.empty((png.height, png.width, 4), np.uint8)
    s = png.as_struct()
    rc = lib().oracle_expand_rgba8(C.byref(s), mode, _u8(out))
    if rc != OK:
        raise ValueError(f"oracle_expand_rgba8 rc={rc}")
    return out


def rgb8(png: DecodedPng):
    out = np.empty((png.height, png.width, 3), np.uint8)
    s = png.as_struct()
    rc = lib().oracle_rgb8(C.byref(s), _u8(out))
    if rc != OK:
        raise ValueError(f"oracle_rgb8 rc={rc}")
    return out


def filter_scanlines(pixels, rows, row_bytes, bpp, paeth_compat=0):
    pixels = np.ascontiguousarray(pixels, np.uint8).reshape(-1)
    stream, ft = np.empty(rows * (row_bytes + 1), np.uint8), np.empty(rows, np.uint8)
    lib().oracle_filter_scanlines(_u8(pixels), rows, row_bytes, bpp, paeth_compat, _u8(stream), _u8(ft))
    return stream, ft


def adam7_pass_sizes(w, h):
    pw, ph = (C.c_uint32 * 7)(), (C.c_uint32 * 7)()
    lib().oracle_adam7_pass_sizes(w, h, pw, ph)
    return list(pw), list(ph)


def front_lut(cs, gamma):
    lut = np.empty(256, np.uint8)
    lib().oracle_front_lut(cs, gamma, _u8(lut))
    return lut


def _img(a):
    a = np.ascontiguousarray(a, np.uint8)
    assert a.ndim == 3 and a.shape[2] == 4
    return a


def front_s(img, cs=CS_REFERENCE, gamma=0, grayscale=0, invert=0):
    img = _img(img)
    out = np.empty_like(img)
    lut = front_lut(cs, gamma)
    lib().oracle_front_s(_u8(img), img.shape[1], img.shape[0], _u8(lut), grayscale, invert, _u8(out))
    return out


def front_r(img, cs=CS_REFERENCE, gamma=0, grayscale=0, invert=0):
    img = _img(img)
    out = np.empty_like(img)
    lib().oracle_front_r(_u8(img), img.shape[1], img.shape[0], cs, gamma, grayscale, invert, _u8(out))
    return out


def blur_taps(radius):
    taps = np.zeros(radius, np.float32)
    dmin = C.c_int()
    rc = lib().oracle_blur_taps(radius, taps.ctypes.data_as(C.POINTER(C.c_float)), C.byref(dmin))
    if rc != OK:
        raise ValueError("bad radius")
    return taps, dmin.value


def _stage(fn, img, *args):
    img = _img(img)
    out = np.empty_like(img)
    rc = fn(_u8(img), img.shape[1], img.shape[0], *args, _u8(out))
    if rc not in (None, OK):
        raise ValueError(f"oracle rc={rc}")
    return out


def blur_s(img, radius):
    return _stage(lib().oracle_blur_s, img, radius)


def blur_r(img, radius):
    return _stage(lib().oracle_blur_r, img, radius)


def sobel_s(img):
    return _stage(lib().oracle_sobel_s, img)


def sobel_r(img):
    return _stage(lib().oracle_sobel_r, img)


def sharpen_s(img, factor):
    return _stage(lib().oracle_sharpen_s, img, factor)


def sharpen_r(img, factor):
    return _stage(lib().oracle_sharpen_r, img, factor)


def box_blur_s(img):
    return _stage(lib().oracle_box_blur_s, img)


def box_blur_r(img):
    return _stage(lib().oracle_box_blur_r, img)


def srgb_tables():
    e, t = np.zeros(256, np.float32), np.zeros(256, np.float32)
    lib().oracle_srgb_tables(e.ctypes.data_as(C.POINTER(C.c_float)), t.ctypes.data_as(C.POINTER(C.c_float)))
    return e, t


def srgb_encode_lut():
    lut = np.empty(256, np.uint8)
    lib().oracle_srgb_encode_lut(_u8(lut))
    return lut


def encode_srgb(img):
    """The surface's linear -> sRGB encode of a finished RGBA8 frame (or batch of frames)."""
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty_like(img)
    lib().oracle_encode_srgb(_u8(img), img.size // 4, _u8(out))
    return out


def overlay(img, circles=(), selected=None, crosshair=None):
    """circles: (x, y, radius) in normalised space; crosshair: (cx, cy, radius) in pixels or None."""
    img = _img(img)
    out = np.empty_like(img)
    su = ShapeUniform(img.shape[1], img.shape[0], len(circles), 0xFFFFFFFF if selected is None else selected)
    arr = (CircleData * max(len(circles), 1))(*[CircleData(x, y, r, 0.0) for x, y, r in circles])
    du = DrawUniform()
    if crosshair is not None:
        du.crosshair, du.circle_center_x, du.circle_center_y, du.circle_radius = 1, *crosshair
    rc = lib().oracle_overlay(_u8(img), img.shape[1], img.shape[0], C.byref(su), arr, C.byref(du), _u8(out))
    if rc != OK:
        raise ValueError(f"oracle rc={rc}")
    return out


def ssim_moments(a, b):
    a, b = _img(a), _img(b)
    assert a.shape == b.shape
    m = np.zeros(5, np.float64)
    lib().oracle_ssim_moments(_u8(a), _u8(b), a.shape[0] * a.shape[1], m.ctypes.data_as(C.POINTER(C.c_double)))
    return m


def ssim_from_moments(m, n):
    m = np.ascontiguousarray(m, np.float64)
    return lib().oracle_ssim_from_moments(m.ctypes.data_as(C.POINTER(C.c_double)), float(n))


def ssim(a, b):
    a, b = _img(a), _img(b)
    assert a.shape == b.shape
    return lib().oracle_ssim(_u8(a), _u8(b), a.shape[0] * a.shape[1])


def ssim_luma(a, b, mode_a, mode_b):
    a, b = _img(a), _img(b)
    assert a.shape == b.shape
    return lib().oracle_ssim_luma(_u8(a), _u8(b), a.shape[0] * a.shape[1], mode_a, mode_b)


def ssim_f32_literal(a, b):
    a, b = _img(a), _img(b)
    return float(lib().oracle_ssim_f32_literal(_u8(a), _u8(b), a.shape[0] * a.shape[1]))


def effects(img, uniform, cs=CS_REFERENCE):
    img = _img(img)
    out = np.empty_like(img)
    rc = lib().oracle_effects(C.byref(uniform), cs, _u8(img), _u8(out))
    if rc != OK:
        raise ValueError(f"oracle_effects rc={rc}")
    return out


def chain(img, uniform, cs=CS_REFERENCE):
    img = _img(img)
    out = np.empty_like(img)
    s = C.c_double()
    rc = lib().oracle_chain(C.byref(uniform), cs, _u8(img), _u8(out), C.byref(s))
    if rc != OK:
        raise ValueError(f"oracle_chain rc={rc}")
    return out, s.value


def chain_batch(frames, uniform, cs=CS_RAW):
    frames = np.ascontiguousarray(frames, np.uint8)
    assert frames.ndim == 4 and frames.shape[3] == 4
    out = np.empty_like(frames)
    s = np.zeros(frames.shape[0], np.float64)
    rc = lib().oracle_chain_batch(C.byref(uniform), cs, _u8(frames), _u8(out),
                                  s.ctypes.data_as(C.POINTER(C.c_double)), frames.shape[0])
    if rc != OK:
        raise ValueError(f"oracle_chain_batch rc={rc}")
    return out, s
