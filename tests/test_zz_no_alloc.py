"""norm_cuda_ssim_static: the SSIM entry point for callers that must not allocate.

Who needs it: a frame loop that has to stay away from the device allocator while work is in
flight (cudaMalloc / cudaFree stall the whole GPU), and a host that already owns all device
memory through its own pool, so that an allocation made by this library would fail.  The entry
point keeps its partial sums in a fixed array inside the library instead of cudaMalloc'd scratch.

This file is collected last on purpose: if the static-scratch path ever breaks, a `pytest -x` run
has already executed every other test.  It is self-contained and also passes in a fresh process.

The check: same images through the allocating entry points (norm_cuda_ssim / _moments) and
through the static one, called twice in a row with different inputs and sizes - the second call
must not see anything of the first in the shared array.  Tolerances are the ones the SSIM parity
tests use (1e-9 absolute on SSIM, 1e-13 relative on the raw moments).
"""
import ctypes as C

import numpy as np
import pytest

import norm_b200 as nb
from conftest import smooth_rgba, splitmix_rgba

pytestmark = pytest.mark.gpu

SSIM_TIGHT = 1e-9


def _static_ssim(lib, a, b, stream, torch):
    """Calls norm_cuda_ssim_static on two DeviceImages; outputs live in caller-owned device memory."""
    n = a.frames
    ssim_dev = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
    mom_dev = torch.full((n, 5), float("nan"), dtype=torch.float64, device="cuda")
    st = lib.norm_cuda_ssim_static(C.c_void_p(a.device_ptr), C.c_void_p(b.device_ptr), a.width, a.height, n,
                                   a.pitch, a.frame_stride, C.c_void_p(ssim_dev.data_ptr()),
                                   C.c_void_p(mom_dev.data_ptr()), C.c_void_p(stream))
    assert st == nb._lib.NORM_OK, lib.norm_cuda_last_error(None).decode()
    torch.cuda.synchronize()
    return ssim_dev.cpu().numpy(), mom_dev.cpu().numpy()


def test_static_scratch_ssim_matches_allocating_entry_point():
    import torch
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    ctx = nb.Context(0, stream=stream)  # the existing entry points, on the same stream
    lib = nb.load()

    # call 1: three ragged 333x200 frames; call 2: one 1920x1080 frame (more CTAs, different
    # slot count per frame) - run back to back so the second finds the first's leftovers
    f1a = np.stack([smooth_rgba(333, 200, seed=s) for s in (1, 2, 3)])
    f1b = np.stack([splitmix_rgba(333, 200, seed=s) for s in (4, 5, 6)])
    f2a = smooth_rgba(1920, 1080, seed=7)[None]
    f2b = (f2a // 2 + splitmix_rgba(1920, 1080, seed=8)[None] // 2).astype(np.uint8)
    imgs = [(ctx.image_from_host(f1a), ctx.image_from_host(f1b)), (ctx.image_from_host(f2a), ctx.image_from_host(f2b))]

    expected = [(ctx.ssim(a, b), ctx.ssim_moments(a, b)) for a, b in imgs]
    for rep in range(2):  # the pair of calls, twice
        got = [_static_ssim(lib, a, b, stream, torch) for a, b in imgs]
        for (gs, gm), (es, em) in zip(got, expected):
            assert np.isfinite(gs).all() and np.abs(gs - es).max() <= SSIM_TIGHT, (rep, gs, es)
            assert np.allclose(gm, em, rtol=1e-13, atol=0), (rep, gm, em)

    # only one of the two outputs requested
    a, b = imgs[0]
    only = torch.zeros(3, dtype=torch.float64, device="cuda")
    st = lib.norm_cuda_ssim_static(C.c_void_p(a.device_ptr), C.c_void_p(b.device_ptr), a.width, a.height, 3, a.pitch,
                                   a.frame_stride, C.c_void_p(only.data_ptr()), None, C.c_void_p(stream))
    assert st == nb._lib.NORM_OK
    torch.cuda.synchronize()
    assert np.abs(only.cpu().numpy() - expected[0][0]).max() <= SSIM_TIGHT

    for a, b in imgs:
        a.destroy(), b.destroy()
    ctx.close()


def test_static_scratch_rejects_oversized_batches_and_bad_arguments():
    import torch
    torch.cuda.set_device(0)
    lib = nb.load()
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    out = torch.zeros(8, dtype=torch.float64, device="cuda")
    p, o = C.c_void_p(buf.data_ptr()), C.c_void_p(out.data_ptr())
    too_many = 1025  # NORM_SSIM_STATIC_MAX_FRAMES + 1
    st = lib.norm_cuda_ssim_static(p, p, 4, 4, too_many, 16, 64, o, None, None)
    assert st == nb._lib.NORM_E_INVALID_ARG
    msg = lib.norm_cuda_last_error(None).decode()
    assert "1025" in msg and "at most 1024 frames" in msg
    assert lib.norm_cuda_ssim_static(p, p, 4, 4, 1, 16, 64, None, None, None) == nb._lib.NORM_E_INVALID_ARG  # no output
    assert lib.norm_cuda_ssim_static(p, p, 4, 4, 1, 12, 64, o, None, None) == nb._lib.NORM_E_INVALID_ARG     # pitch
    assert lib.norm_cuda_ssim_static(None, p, 4, 4, 1, 16, 64, o, None, None) == nb._lib.NORM_E_INVALID_ARG  # NULL image
    # and a valid call still works afterwards
    assert lib.norm_cuda_ssim_static(p, p, 4, 4, 1, 16, 64, o, None, None) == nb._lib.NORM_OK
    torch.cuda.synchronize()
    assert abs(float(out[0].item()) - 1.0) <= 1e-12  # identical images
