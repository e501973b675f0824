"""ctypes view of libnorm_nccl.so (include/norm_nccl.h): the SSIM-moments all-reduce."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnorm_nccl.so")

PROTOTYPES = {
    "norm_nccl_unique_id": (C.c_int32, [C.c_void_p]),
    "norm_nccl_comm_create": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "norm_nccl_comm_destroy": (None, [C.c_void_p]),
    "norm_nccl_allreduce_sum_f64": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "norm_nccl_allreduce_moments": (C.c_int32, [C.c_void_p, C.POINTER(C.c_double), C.c_size_t]),
    "norm_nccl_version": (C.c_int32, []),
    "norm_nccl_last_error": (C.c_char_p, []),
}
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing; build it with `python build.py`")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def unique_id():
    buf = (C.c_uint8 * 128)()
    if load().norm_nccl_unique_id(buf):
        raise RuntimeError(load().norm_nccl_last_error().decode())
    return bytes(buf)


class Comm:
    """One rank's NCCL communicator on the current CUDA device."""

    def __init__(self, uid: bytes, nranks: int, rank: int):
        self.lib = load()
        h = C.c_void_p()
        buf = (C.c_uint8 * 128).from_buffer_copy(uid)
        if self.lib.norm_nccl_comm_create(buf, nranks, rank, C.byref(h)):
            raise RuntimeError(self.lib.norm_nccl_last_error().decode())
        self.h = h

    def allreduce_moments(self, m):
        m = np.ascontiguousarray(m, np.float64).copy()
        if self.lib.norm_nccl_allreduce_moments(self.h, m.ctypes.data_as(C.POINTER(C.c_double)), m.size):
            raise RuntimeError(self.lib.norm_nccl_last_error().decode())
        return m

    def allreduce_device(self, dev_ptr, count, stream=0):
        if self.lib.norm_nccl_allreduce_sum_f64(self.h, C.c_void_p(dev_ptr), count, C.c_void_p(stream)):
            raise RuntimeError(self.lib.norm_nccl_last_error().decode())

    def close(self):
        if self.h:
            self.lib.norm_nccl_comm_destroy(self.h)
            self.h = None
