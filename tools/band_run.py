#!/usr/bin/env python
"""BASELINE configs[4]: ONE giant RGBA8 image, row-band sharded over the GPUs of one box.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/band_run.py --size 32768 [--radius 21] [--verify]

One process per GPU.  Each rank owns H/N rows, pulls the stencil halo rows (11 per side at r=21
with Sobel) from its neighbours' device memory with peer copies over NVLink (CUDA IPC mappings,
norm_cuda_ipc_* + norm_cuda_copy_rows), runs the fused chain on band + halo
(norm_cuda_chain_band) and sums the five SSIM moments of its owned rows over ranks with NCCL
(libnorm_nccl.so).  torch.distributed (gloo) is only the host side-channel for the handles.
Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def gen_rows(xp, y0, y1, w, device=None):
    """Deterministic synthetic rows [y0,y1) x w x 4 (same integers from numpy and torch)."""
    if xp is np:
        y = np.arange(y0, y1, dtype=np.int64)[:, None, None]
        x = np.arange(w, dtype=np.int64)[None, :, None]
        c = np.arange(4, dtype=np.int64)[None, None, :]
    else:
        y = xp.arange(y0, y1, dtype=xp.int64, device=device)[:, None, None]
        x = xp.arange(w, dtype=xp.int64, device=device)[None, :, None]
        c = xp.arange(4, dtype=xp.int64, device=device)[None, None, :]
    h = (x * 73856093) ^ (y * 19349663) ^ (c * 83492791)
    h = (h ^ (h >> 13)) * 1274126177
    noise = (h >> 7) & 127
    smooth = ((x >> 3) * 5 + (y >> 2) * 3 + c * 40) & 127
    v = (noise + smooth) & 255
    if xp is np:
        v = v.astype(np.uint8)
        v[..., 3] = 255
    else:
        v = v.to(xp.uint8)
        v[..., 3] = 255
    return v


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=32768)
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--radius", type=int, default=21)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--verify", action="store_true")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import norm_b200 as nb
    from norm_b200 import nccl

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, W = a.size, (a.width or a.size)

    ctx = nb.Context(local)
    f = nb.FeatureUniform(W, H)
    f.raw.blur, f.raw.blur_radius, f.raw.edge_detect = 1, a.radius, 1
    lay = nb.sharding.band_layout(f, H, world, rank)
    src = ctx.image_create(W, lay["ext_rows"], 1)
    dst = ctx.image_create(W, lay["ext_rows"], 1)
    # own rows generated on the device, in slabs
    slab = max(1, (256 << 20) // (W * 4))
    for y in range(lay["row0"], lay["row0"] + lay["rows"], slab):
        y1 = min(y + slab, lay["row0"] + lay["rows"])
        t = gen_rows(torch, y, y1, W, device=f"cuda:{local}").contiguous()
        torch.cuda.synchronize()  # torch's stream and the library's stream are independent
        ctx.copy_rows(src, lay["own_off"] + (y - lay["row0"]), t.data_ptr(), W * 4, 0, y1 - y)
        ctx.sync()
        del t
    # exchange IPC handles, map the neighbours
    handles = [None] * world
    dist.all_gather_object(handles, (ctx.ipc_export(src), src.pitch))
    reqs = nb.sharding.halo_requests(f, H, world, rank)
    peers = {p: ctx.ipc_open(handles[p][0]) for p in {r[0] for r in reqs}}
    uid = [nccl.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    comm = nccl.Comm(uid[0], world, rank)
    fb = nb.FeatureUniform(W, lay["ext_rows"])
    fb.raw.blur, fb.raw.blur_radius, fb.raw.edge_detect = 1, a.radius, 1
    dist.barrier()

    # everything of a step is queued on the context's stream: peer halo copies, the fused chain, and the
    # in-place NCCL all-reduce of the 5 moments in device memory; the host waits once per step
    mom = ctx.image_create(64, 1, 1)  # 256 bytes of device memory for the moments
    stream = ctx.stream()

    def step():
        for (p, prow, mine, n) in reqs:  # halo rows over NVLink
            ctx.copy_rows(src, mine, peers[p], handles[p][1], prow, n)
        ctx.chain_band_async(fb, src, dst, lay["own_off"], lay["rows"], nb.NORM_CS_RAW, mom.device_ptr)
        comm.allreduce_device(mom.device_ptr, 5, stream)
        ctx.sync()

    for _ in range(a.warmup):
        step()
    dist.barrier()
    t0 = time.perf_counter()
    ctx.timer_start()
    for _ in range(a.steps):
        step()
    ctx.timer_stop()
    dev_ms = ctx.timer_elapsed_ms() / a.steps
    wall_ms = (time.perf_counter() - t0) / a.steps * 1e3
    tot = np.frombuffer(mom.download()[0].tobytes()[:40], np.float64).copy()
    tmax = torch.tensor([dev_ms, wall_ms], dtype=torch.float64)
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ssim = nb.ssim_from_moments(tot, float(W) * float(H))

    ok = None
    if a.verify:
        import oracle_lib as oracle
        ok = True
        # every seam this rank touches: rows around its band edges against the CPU oracle on a strip
        for seam in {lay["row0"], lay["row0"] + lay["rows"]} - {0, H}:
            lo, hi = max(0, seam - 40), min(H, seam + 40)
            strip = gen_rows(np, lo, hi, W)
            eo, _ = oracle.chain(strip, oracle.make_uniform(W, hi - lo, blur=1, blur_radius=a.radius, edge_detect=1),
                                 oracle.CS_RAW)
            # rows of the strip far enough from its own artificial edges, and owned by this rank
            g0, g1 = max(lo + 14, lay["row0"]), min(hi - 14, lay["row0"] + lay["rows"])
            if lo == 0:
                g0 = max(0, lay["row0"])
            if hi == H:
                g1 = min(H, lay["row0"] + lay["rows"])
            full = dst.download()[0] if lay["ext_rows"] * W * 4 < (1 << 31) else None
            if full is None:
                tmp = ctx.image_create(W, g1 - g0, 1)
                ctx.copy_rows(tmp, 0, dst.device_ptr, dst.pitch, lay["own_off"] + (g0 - lay["row0"]), g1 - g0)
                ctx.sync()
                got = tmp.download()[0]
                tmp.destroy()
            else:
                got = full[lay["own_off"] + (g0 - lay["row0"]): lay["own_off"] + (g1 - lay["row0"])]
            same = bool(np.array_equal(got, eo[g0 - lo:g1 - lo]))
            if not same and os.environ.get("BAND_DEBUG"):
                bad = np.where((got != eo[g0 - lo:g1 - lo]).any(axis=(1, 2)))[0]
                srow = src.download()[0][lay["own_off"] + (g0 - lay["row0"]): lay["own_off"] + (g1 - lay["row0"])]
                print(f"rank {rank} seam {seam}: rows {g0}..{g1} bad rows {bad[:8] + g0} n={bad.size} "
                      f"src_ok={np.array_equal(srow, strip[g0 - lo:g1 - lo])}", flush=True)
                halo = src.download()[0]
                for (p_, prow, mine, n) in reqs:
                    gy = lay["ext_row0"] + mine
                    print(f"  halo from {p_}: local {mine}+{n} global {gy}: ok={np.array_equal(halo[mine:mine + n], gen_rows(np, gy, gy + n, W))}", flush=True)
            ok = ok and same
        oks = [None] * world
        dist.all_gather_object(oks, ok)
        ok = all(oks)
    if rank == 0:
        px = float(W) * float(H)
        print(json.dumps({"workload": f"single {W}x{H} RGBA8 image, row-band sharded, blur r={a.radius} + Sobel + SSIM",
                          "n_gpus": world, "ms_per_step_device_max": float(tmax[0]), "ms_per_step_wall_max": float(tmax[1]),
                          "Mpixel_s": px / (float(tmax[1]) * 1e-3) / 1e6, "ssim": ssim,
                          "halo_rows_per_side": lay["down"] or lay["up"], "seams_bit_exact_vs_oracle": ok,
                          "nccl_version": nccl.load().norm_nccl_version()}))
    for p in peers.values():
        ctx.ipc_close(p)
    comm.close()
    src.destroy(), dst.destroy(), mom.destroy()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
