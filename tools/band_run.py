#!/usr/bin/env python
"""BASELINE configs[4]: ONE giant RGBA8 image, row-band sharded over the GPUs of one box.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/band_run.py --size 32768 [--radius 21] [--verify full|seams|none]

One process per GPU.  Each rank owns H/N rows; norm_cuda_band_* (include/norm_cuda.h) pulls the stencil halo rows
(11 per side at r=21 with Sobel) out of the neighbours' device memory with peer copies over NVLink on a halo
stream, runs the fused chain on the interior rows meanwhile and on the two edge strips afterwards, and hands the
five SSIM moments of the owned rows to an NCCL all-reduce (libnorm_nccl.so) on a side stream.  No host
synchronisation inside the step loop.  `band_benchmark` is also what bench.py's "band" key runs.
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


def band_benchmark(torch, dist, ctx, local_rank, rank, world, height, width, radius, steps, warmup, verify="full",
                   group=None, use_nccl=True):
    """Runs the band-sharded chain (blur r + Sobel + SSIM, colour space raw) on a synthetic height x width image.
    dist: torch.distributed (any backend; used for the handle exchange, the enqueue barrier and the max over ranks).
    -> dict for rank 0's JSON line (every rank returns it)."""
    import norm_b200 as nb
    from norm_b200 import nccl
    from synth import gen_rows

    dev = f"cuda:{local_rank}"
    f = nb.FeatureUniform(width, height)
    f.raw.blur, f.raw.blur_radius, f.raw.edge_detect = 1, radius, 1
    f_kw = dict(blur=1, blur_radius=radius, edge_detect=1)

    def allgather(obj):
        out = [None] * world
        dist.all_gather_object(out, obj, group=group)
        return out

    def barrier():
        # with the NCCL backend dist.barrier() only queues a kernel: wait for it, or the ranks enter the timed loop
        # skewed and the wall clock of the early ones contains the wait for the late ones
        torch.cuda.synchronize()
        dist.barrier(group=group)
        torch.cuda.synchronize()

    comm = None
    if use_nccl and world > 1:
        uid = [nccl.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0, group=group)
        comm = nccl.Comm(uid[0], world, rank)
    job = nb.sharding.BandJob(ctx, f, width, height, world, rank, allgather, comm)
    b = job.band
    # own rows generated on the device, in slabs
    slab = max(1, (256 << 20) // (width * 4))
    for y in range(0, b.rows, slab):
        n = min(slab, b.rows - y)
        t = gen_rows(torch, b.row0 + y, b.row0 + y + n, width, device=dev).contiguous()
        torch.cuda.synchronize()  # torch's stream and the library's stream are independent
        job.fill_rows(t.data_ptr(), width * 4, y, n)
        ctx.sync()
        del t
    job.publish()
    ctx.sync()
    barrier()  # every rank has published its rows

    for _ in range(max(warmup, 1)):
        job.step(nb.NORM_CS_RAW)
    job.sync()
    barrier()
    t0 = time.perf_counter()
    ctx.timer_start()
    for _ in range(steps):
        job.step(nb.NORM_CS_RAW)
    ctx.timer_stop()
    job.sync()
    wall_ms = (time.perf_counter() - t0) / steps * 1e3
    dev_ms = ctx.timer_elapsed_ms() / steps
    tmax = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device=dev if dist.get_backend(group) == "nccl" else "cpu")
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX, group=group)
    tot = job.moments()  # summed over ranks by the all-reduce (or this rank's alone without a communicator)
    ssim = nb.ssim_from_moments(tot, float(width) * float(height))

    # the moments all-reduce alone: 5 doubles, latency-bound
    ar_us = None
    if comm is not None:
        stream = job.side
        for _ in range(5):
            comm.allreduce_device(job.mom[0].device_ptr, 5, stream)
        job.sync()
        barrier()
        t1 = time.perf_counter()
        for _ in range(100):
            comm.allreduce_device(job.mom[0].device_ptr, 5, stream)
        job.sync()
        ar_us = (time.perf_counter() - t1) / 100 * 1e6

    res = {"workload": f"single {width}x{height} RGBA8 image, row-band sharded over {world} GPUs, blur r={radius} + Sobel + SSIM",
           "n_gpus": world, "ms_per_step": float(tmax[1]), "ms_per_step_device_main_stream": float(tmax[0]),
           "Mpixel_s": float(width) * float(height) / (float(tmax[1]) * 1e-3) / 1e6,
           "halo_rows_per_side": max(b.halo_up, b.halo_down), "halo_bytes_per_step_per_gpu": job.halo_bytes,
           "ssim": ssim, "allreduce_us": ar_us, "overlap": "interior rows run while the halo rows are copied; "
           "all-reduce on a side stream; no host sync inside the step loop"}
    if verify != "none":
        import oracle_lib as oracle
        from band_check import check_band
        oracle.build()
        oracle.lib().oracle_set_num_threads(max(1, len(os.sched_getaffinity(0)) // world))
        chk = check_band(job, f_kw, width, height, full=(verify == "full"))
        flags = torch.tensor([0 if chk["rows_bit_exact"] is False else 1, 1 if chk["seams_bit_exact"] else 0,
                              1 if chk["halo_rows_ok"] else 0], dtype=torch.float64, device=tmax.device)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN, group=group)
        res["seams_bit_exact_vs_oracle"] = bool(flags[1] > 0)
        res["halo_rows_equal_generator"] = bool(flags[2] > 0)
        if verify == "full":
            res["full_image_bit_exact_vs_oracle"] = bool(flags[0] > 0)
            om = torch.tensor(chk["oracle_moments"], dtype=torch.float64, device=tmax.device)
            dist.all_reduce(om, op=dist.ReduceOp.SUM, group=group)
            res["ssim_abs_err_vs_oracle"] = abs(ssim - nb.ssim_from_moments(om.cpu().numpy(), float(width) * float(height)))
    job.close()
    if comm is not None:
        res["nccl_version"] = nccl.load().norm_nccl_version()
        comm.close()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=32768)
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--radius", type=int, default=21)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--verify", default="seams", choices=["full", "seams", "none"])
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--no-nccl", action="store_true", help="skip the all-reduce (timing experiments)")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import norm_b200 as nb

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ctx = nb.Context(local)
    res = band_benchmark(torch, dist, ctx, local, rank, world, a.height or a.size, a.width or a.size, a.radius, a.steps, a.warmup, a.verify,
                         use_nccl=not a.no_nccl)
    if rank == 0:
        print(json.dumps(res))
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
