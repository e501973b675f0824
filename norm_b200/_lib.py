"""ctypes binding of norm_b200/libnorm_cuda.so (C ABI: include/norm_cuda.h).

The library is the product; this module only loads it and declares prototypes.  There is no
fallback: if the shared object is missing, loading raises, and if no CUDA device is present
`norm_cuda_create` fails with NORM_E_CUDA.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnorm_cuda.so")

NORM_OK, NORM_E_INVALID_ARG, NORM_E_UNSUPPORTED, NORM_E_CUDA, NORM_E_NCCL, NORM_E_OOM, NORM_E_DIM_MISMATCH = range(7)
NORM_CS_REFERENCE, NORM_CS_RAW = 0, 1
NORM_OUT_SRGB = 2
NORM_LUMA_RGB, NORM_LUMA_GRAY, NORM_LUMA_GRAY_UNIT = 0, 1, 2
NORM_EXPAND_COMPAT, NORM_EXPAND_SPEC = 0, 1
NORM_MAX_BLUR_RADIUS = 63


class NormFeatureUniform(C.Structure):
    """src/renderer/feature_uniform.rs:1-17 -- 112 bytes, repr(C)."""
    _fields_ = [
        ("grayscale", C.c_uint32), ("invert", C.c_uint32), ("gamma", C.c_uint32),
        ("blur", C.c_uint32), ("blur_radius", C.c_uint32), ("width", C.c_uint32),
        ("height", C.c_uint32), ("sharpen", C.c_uint32), ("sharpen_factor", C.c_uint32),
        ("edge_detect", C.c_uint32), ("_padding_1", C.c_uint8 * 8), ("transform", (C.c_float * 4) * 4),
    ]


assert C.sizeof(NormFeatureUniform) == 112


class NormInflatedDesc(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("bit_depth", C.c_uint8), ("color_type", C.c_uint8), ("interlace", C.c_uint8), ("_reserved", C.c_uint8),
        ("data", C.c_void_p), ("len", C.c_size_t),
    ]


class NormPngDesc(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("bit_depth", C.c_uint8), ("color_type", C.c_uint8), ("_reserved", C.c_uint8 * 2),
        ("gamma", C.c_uint32),
        ("pixels", C.c_void_p), ("len", C.c_size_t),
        ("plte", C.c_void_p), ("n_plte", C.c_int32),
        ("trns", C.c_void_p), ("n_trns", C.c_int32),
    ]


class NormShapeUniform(C.Structure):
    """src/renderer/shape_uniform.rs:3-10"""
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("num_circles", C.c_uint32),
                ("selected_circle", C.c_uint32)]


class NormCircleData(C.Structure):
    """src/renderer/shape_uniform.rs:38-45"""
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("radius", C.c_float), ("_padding", C.c_float)]


class NormDrawUniform(C.Structure):
    """src/renderer/draw_uniform.rs:3-9"""
    _fields_ = [("crosshair", C.c_uint32), ("circle_center_x", C.c_float), ("circle_center_y", C.c_float),
                ("circle_radius", C.c_float), ("camera_view_proj", (C.c_float * 4) * 4)]


class NormBandHandle(C.Structure):
    """include/norm_cuda.h: what one rank of a row-band sharded image tells the others (256 bytes)."""
    _fields_ = [("mem", C.c_uint8 * 64), ("ready_event", C.c_uint8 * 64), ("halo_event", C.c_uint8 * 64),
                ("pitch", C.c_uint64), ("local_ptr", C.c_uint64), ("local_ready", C.c_uint64),
                ("local_halo", C.c_uint64), ("pid", C.c_uint64),
                ("row0", C.c_uint32), ("rows", C.c_uint32), ("halo_up", C.c_uint32), ("halo_down", C.c_uint32),
                ("rank", C.c_int32), ("device", C.c_int32)]


assert C.sizeof(NormBandHandle) == 256
assert C.sizeof(NormShapeUniform) == 16 and C.sizeof(NormCircleData) == 16 and C.sizeof(NormDrawUniform) == 80
NORM_MAX_CIRCLES = 256

_vp, _u32, _i32, _sz = C.c_void_p, C.c_uint32, C.c_int32, C.c_size_t
_fu = C.POINTER(NormFeatureUniform)
_dp = C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol include/norm_cuda.h declares
PROTOTYPES = {
    "norm_cuda_abi_version": (_i32, []),
    "norm_cuda_status_string": (C.c_char_p, [_i32]),
    "norm_cuda_create": (_i32, [_i32, C.POINTER(_vp)]),
    "norm_cuda_create_on_stream": (_i32, [_i32, _vp, C.POINTER(_vp)]),
    "norm_cuda_destroy": (None, [_vp]),
    "norm_cuda_last_error": (C.c_char_p, [_vp]),
    "norm_cuda_sync": (_i32, [_vp]),
    "norm_cuda_launch_count": (C.c_uint64, [_vp]),
    "norm_cuda_device_info": (_i32, [_vp, C.POINTER(_i32), C.POINTER(_sz), C.c_char_p, _sz]),
    "norm_cuda_timer_start": (_i32, [_vp]),
    "norm_cuda_timer_stop": (_i32, [_vp]),
    "norm_cuda_timer_elapsed_ms": (_i32, [_vp, C.POINTER(C.c_float)]),
    "norm_cuda_image_create": (_i32, [_vp, _u32, _u32, _u32, C.POINTER(_vp)]),
    "norm_cuda_image_wrap": (_i32, [_vp, _vp, _u32, _u32, _u32, _sz, _sz, C.POINTER(_vp)]),
    "norm_cuda_image_destroy": (None, [_vp, _vp]),
    "norm_cuda_image_info": (_i32, [_vp, C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_sz),
                                    C.POINTER(_sz), C.POINTER(_vp)]),
    "norm_cuda_upload": (_i32, [_vp, _vp, _u32, _u32, _vp, _sz]),
    "norm_cuda_download": (_i32, [_vp, _vp, _u32, _u32, _vp, _sz]),
    "norm_cuda_host_alloc": (_i32, [_vp, _sz, C.POINTER(_vp)]),
    "norm_cuda_host_free": (None, [_vp, _vp]),
    "norm_cuda_expand_rgba8": (_i32, [_vp, C.POINTER(NormPngDesc), _i32, _vp, _u32]),
    "norm_cuda_expand_rgba8_host": (_i32, [_vp, C.POINTER(NormPngDesc), _i32, _vp]),
    "norm_cuda_expand_rgba8_device": (_i32, [_vp, C.POINTER(NormPngDesc), _sz, _sz, _u32, _i32, _vp, _u32]),
    "norm_cuda_unfilter": (_i32, [_vp, C.POINTER(NormInflatedDesc), _vp]),
    "norm_cuda_unfilter_expand_rgba8": (_i32, [_vp, C.POINTER(NormInflatedDesc), _vp, _i32, _vp, _i32, _i32, _vp, _u32]),
    "norm_cuda_unfilter_expand_rgba8_batch": (_i32, [_vp, C.POINTER(NormInflatedDesc), _u32, _vp, _i32, _vp, _i32, _i32,
                                                     _vp, _u32]),
    "norm_cuda_filter_scanlines": (_i32, [_vp, _u32, _u32, C.c_uint8, C.c_uint8, _vp, _i32, _vp, _vp]),
    "norm_cuda_filter_image": (_i32, [_vp, _vp, _u32, _i32, _vp, _vp]),
    "norm_cuda_effects": (_i32, [_vp, _fu, _u32, _vp, _vp]),
    "norm_cuda_chain": (_i32, [_vp, _fu, _u32, _vp, _vp, _dp]),
    "norm_cuda_chain_async": (_i32, [_vp, _fu, _u32, _vp, _vp, _vp]),
    "norm_cuda_chain_host": (_i32, [_vp, _fu, _u32, _vp, _vp, _u32, _dp]),
    "norm_cuda_chain_band": (_i32, [_vp, _fu, _u32, _vp, _vp, _u32, _u32, _dp]),
    "norm_cuda_chain_band_async": (_i32, [_vp, _fu, _u32, _vp, _vp, _u32, _u32, _vp]),
    "norm_cuda_stream": (_i32, [_vp, C.POINTER(_vp)]),
    "norm_cuda_ipc_export": (_i32, [_vp, _vp, _vp]),
    "norm_cuda_ipc_open": (_i32, [_vp, _vp, C.POINTER(_vp)]),
    "norm_cuda_ipc_close": (_i32, [_vp, _vp]),
    "norm_cuda_copy_rows": (_i32, [_vp, _vp, _u32, _vp, _sz, _u32, _u32]),
    "norm_cuda_band_create": (_i32, [_vp, _fu, _u32, _u32, _i32, _i32, C.POINTER(_vp)]),
    "norm_cuda_band_destroy": (_i32, [_vp]),
    "norm_cuda_band_layout": (_i32, [_vp, C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u32)]),
    "norm_cuda_band_src": (_vp, [_vp]),
    "norm_cuda_band_dst": (_vp, [_vp]),
    "norm_cuda_band_export": (_i32, [_vp, C.POINTER(NormBandHandle)]),
    "norm_cuda_band_connect": (_i32, [_vp, C.POINTER(NormBandHandle), _i32]),
    "norm_cuda_band_begin_update": (_i32, [_vp]),
    "norm_cuda_band_publish": (_i32, [_vp]),
    "norm_cuda_band_step_async": (_i32, [_vp, _u32, _vp]),
    "norm_cuda_band_side_stream": (_i32, [_vp, C.POINTER(_vp)]),
    "norm_cuda_band_sync": (_i32, [_vp]),
    "norm_cuda_front": (_i32, [_vp, _fu, _u32, _vp, _vp]),
    "norm_cuda_blur": (_i32, [_vp, _u32, _vp, _vp]),
    "norm_cuda_sobel": (_i32, [_vp, _vp, _vp]),
    "norm_cuda_sharpen": (_i32, [_vp, _u32, _vp, _vp]),
    "norm_cuda_box_blur": (_i32, [_vp, _vp, _vp]),
    "norm_cuda_overlay": (_i32, [_vp, C.POINTER(NormShapeUniform), C.POINTER(NormCircleData),
                                 C.POINTER(NormDrawUniform), _vp, _vp]),
    "norm_cuda_ssim": (_i32, [_vp, _vp, _vp, _dp]),
    "norm_cuda_ssim_moments": (_i32, [_vp, _vp, _vp, _dp]),
    "norm_cuda_ssim_luma": (_i32, [_vp, _vp, _i32, _vp, _i32, _dp]),
    "norm_ssim_from_moments": (C.c_double, [_dp, C.c_double]),
    "norm_cuda_ssim_static": (_i32, [_vp, _vp, _u32, _u32, _u32, _sz, _sz, _vp, _vp, _vp]),
    "norm_blur_taps": (_i32, [_u32, C.POINTER(C.c_float), C.POINTER(_i32)]),
    "norm_front_lut": (_i32, [_u32, _u32, C.POINTER(C.c_uint8)]),
    "norm_srgb_tables": (_i32, [C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "norm_srgb_encode_lut": (_i32, [C.POINTER(C.c_uint8)]),
    "norm_plan_band": (_i32, [_fu, _u32, _i32, _i32, C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u32),
                              C.POINTER(_u32)]),
    "norm_plan_frames": (_i32, [_u32, _i32, _i32, C.POINTER(_u32), C.POINTER(_u32)]),
}

_lib = None


def load():
    """Loads libnorm_cuda.so (building nothing: run `python build.py` first)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA extension is the only implementation "
                "(no CPU fallback).  Build it with `python build.py`.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib
