"""Host-side logic of the N>1 paths, run on CPU with the gloo backend (world_size 2):
the plans partition the work exactly, a band + halo processed as its own image reproduces the
owned rows of the full-image result, and all-reduced SSIM moments finalise to the full-image
SSIM.  The CPU oracle plays the per-rank compute; libnorm_cuda.so contributes only its
host-only planning / finalising entry points (no GPU here)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import norm_b200 as nb
from conftest import smooth_rgba


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _uniform(oracle, w, h, **kw):
    return oracle.make_uniform(w, h, **kw)


def _band_worker(rank, world, port, w, h, kw, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_lib as oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    oracle.lib().oracle_set_num_threads(2)
    img = smooth_rgba(w, h, seed=17)  # every rank can regenerate any row: stands in for the peer copy
    f = nb.FeatureUniform(w, h)
    for k, v in kw.items():
        setattr(f.raw, k, v)
    lay = nb.sharding.band_layout(f, h, world, rank)
    # the halo plan must name exactly the rows outside the owned range, from their owners
    reqs = nb.sharding.halo_requests(f, h, world, rank)
    got_rows = sorted(r for (_, _, mine, n) in reqs for r in range(mine, mine + n))
    exp_rows = [r for r in range(lay["ext_rows"]) if not lay["own_off"] <= r < lay["own_off"] + lay["rows"]]
    assert got_rows == exp_rows
    for (peer, prow, mine, n) in reqs:
        pl = nb.sharding.band_layout(f, h, world, peer)
        assert pl["ext_row0"] + prow == lay["ext_row0"] + mine  # same global row on both sides
    ext = np.ascontiguousarray(img[lay["ext_row0"]:lay["ext_row0"] + lay["ext_rows"]])
    out_ext, _ = oracle.chain(ext, _uniform(oracle, w, lay["ext_rows"], **kw), oracle.CS_RAW)
    own = slice(lay["own_off"], lay["own_off"] + lay["rows"])
    m = oracle.ssim_moments(np.ascontiguousarray(out_ext[own]), np.ascontiguousarray(ext[own]))
    t = torch.from_numpy(m.copy())
    dist.all_reduce(t)
    ssim = nb.ssim_from_moments(t.numpy(), w * h)
    gathered = [None] * world
    dist.all_gather_object(gathered, (lay["row0"], out_ext[own]))
    if rank == 0:
        full = np.concatenate([g[1] for g in sorted(gathered, key=lambda g: g[0])])
        ref_out, ref_ssim = oracle.chain(img, _uniform(oracle, w, h, **kw), oracle.CS_RAW)
        q.put((bool(np.array_equal(full, ref_out)), float(ssim), float(ref_ssim)))
    dist.destroy_process_group()


@pytest.mark.parametrize("kw", [dict(blur=1, blur_radius=21, edge_detect=1), dict(blur=1, blur_radius=4, sharpen=1),
                                dict(grayscale=1, edge_detect=1)])
def test_band_sharding_two_ranks_gloo(kw):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_band_worker, args=(r, 2, port, 96, 77, kw, q)) for r in range(2)]
    [p.start() for p in procs]
    same, ssim, ref = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert same and abs(ssim - ref) <= 1e-9


def _frames_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_lib as oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    oracle.lib().oracle_set_num_threads(2)
    n, w, h = 5, 64, 40
    frames = np.stack([smooth_rgba(w, h, seed=i) for i in range(n)])
    f0, cnt = nb.plan_frames(n, world, rank)
    u = oracle.make_uniform(w, h, blur=1, blur_radius=7, edge_detect=1)
    _, s = oracle.chain_batch(frames[f0:f0 + cnt], u, oracle.CS_RAW)
    gathered = [None] * world
    dist.all_gather_object(gathered, (f0, s))
    if rank == 0:
        allv = np.concatenate([g[1] for g in sorted(gathered, key=lambda g: g[0])])
        _, ref = oracle.chain_batch(frames, u, oracle.CS_RAW)
        q.put(bool(np.array_equal(allv, ref)))
    dist.destroy_process_group()


def test_frame_sharding_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_frames_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    assert q.get(timeout=120)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)


def test_halo_plan_spans_several_neighbours():
    """Bands shorter than the halo: the halo comes from more than one peer."""
    f = nb.FeatureUniform(64, 40)
    f.toggle_blur(), f.toggle_edge_detect()  # r=21 -> 11 halo rows, bands of 5 rows
    reqs = nb.sharding.halo_requests(f, 40, 8, 3)
    assert {p for p, *_ in reqs} == {0, 1, 2, 4, 5, 6}
    assert sum(n for *_, n in reqs) == 22


def test_nccl_library_exports():
    import ctypes as C
    import re
    from conftest import ROOT
    hdr = open(os.path.join(ROOT, "include", "norm_nccl.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(norm_nccl_[a-z0-9_]+)\s*\(", hdr))
    from norm_b200 import nccl
    lib = C.CDLL(nccl.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(nccl.PROTOTYPES)
    assert nccl.load().norm_nccl_version() >= 22000
