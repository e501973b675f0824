#!/usr/bin/env python
"""Headline benchmark: Mpixel/s of the blur + Sobel + SSIM chain on batched 1080p RGBA8 frames.

    python bench.py [--gpus N] [--steps K] [--warmup W]                       (this repo's CUDA path)
    python bench.py --impl reference [--gpus N] [--steps K] [--warmup W]      (CPU arm)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[3]): ONE batch of 4096 frames of 1920x1080 RGBA8, Gaussian blur at the reference's
default radius 21 -> Sobel line detection -> global SSIM of the result against the source frame.  Frames are
independent, so the batch is sharded over the N ranks with norm_plan_frames and no data-path collective runs
(`"scaling": "strong"`: the total work is fixed, `value` = all 4096 frames' pixels / max-over-ranks device time).

A step = one pass of the chain over the rank's shard, inputs resident in HBM.  Even at N = 8 a shard (4.2 GB) is
~34x the 126 MB L2, so no explicit L2 flush is needed between steps.
After the timed region 12 frames of the TIMED batch (global frames 0, 1, 2047, 4095 + 8 seeded-random ones) are
downloaded with their results and compared with the CPU oracle: bytes equal, SSIM within 1e-9.
`e2e` = the same chain through norm_cuda_chain_host on pinned HOST buffers (H2D + D2H in the timed region) over a
bounded subset of the shard; `e2e.pcie` = the pinned-copy ceilings measured in the same process.
At N > 1 the `band` key runs BASELINE.json configs[4] (one 32768^2 image, row-band sharded, NVLink halo copies,
NCCL all-reduce of the SSIM moments) through norm_cuda_band_* and checks the whole result against the oracle.
The CPU arm / cpu_baseline time oracle/libnorm_oracle.so (a C restatement of the reference; the reference itself
is Rust + WGSL and cannot be built or run here) on all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, H = 1920, 1080
PX = W * H


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=4096, help="frames of the whole batch (sharded over the ranks)")
    ap.add_argument("--band-size", type=int, default=32768, help="side of the band-sharded image of the `band` key (N > 1)")
    ap.add_argument("--no-band", action="store_true")
    ap.add_argument("--radius", type=int, default=21)
    ap.add_argument("--e2e-frames", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true", help="skip the radius 3/7/39 sweep")
    return ap.parse_args()


# --------------------------------------------------------------------------------------- clocks

class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ts, line in self.lines:
            if not (t0 <= ts <= t1 + 0.2):
                continue
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx = float(p[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------- CPU arm

def host_threads(orc):
    """All host cores this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1, so the count is set
    explicitly instead of being inherited."""
    n = len(os.sched_getaffinity(0))
    orc.lib().oracle_set_num_threads(n)
    return n


def cpu_chain_rate(radius, seconds=12.0):
    """Oracle (C port of the reference path) on all host threads, bounded sample. -> dict"""
    import numpy as np
    import oracle_lib as orc
    orc.build()
    nthr = host_threads(orc)
    rng = np.random.default_rng(0)
    n = max(nthr, 8)
    frames = rng.integers(0, 256, size=(n, H, W, 4), dtype=np.uint8)
    f = orc.make_uniform(W, H, blur=1, blur_radius=radius, edge_detect=1)
    orc.chain_batch(frames[:min(n, nthr)], f, orc.CS_RAW)  # warm (page faults, thread pool)
    done, t0 = 0, time.perf_counter()
    while True:
        orc.chain_batch(frames, f, orc.CS_RAW)
        done += n
        dt = time.perf_counter() - t0
        if dt >= seconds:
            break
    out = {"value": done * PX / dt / 1e6, "unit": "Mpixel/s", "cores": nthr, "kind": "port",
           "sample": f"{done} frames 1920x1080 (blur r={radius} + Sobel + SSIM) in {dt:.1f} s, "
                     f"oracle/libnorm_oracle.so with OpenMP over frames, {nthr} threads"}
    # the same port on ONE thread (SURVEY 8d asks for both): a few frames, about 3 s
    orc.lib().oracle_set_num_threads(1)
    done1, t0 = 0, time.perf_counter()
    while True:
        orc.chain_batch(frames[:1], f, orc.CS_RAW)
        done1 += 1
        dt1 = time.perf_counter() - t0
        if dt1 >= 3.0:
            break
    orc.lib().oracle_set_num_threads(nthr)
    out["value_1_thread"] = done1 * PX / dt1 / 1e6
    return out


def run_reference_arm(args, rank):
    if rank != 0:
        return
    import numpy as np
    import oracle_lib as orc
    orc.build()
    nthr = host_threads(orc)
    n = max(nthr, 8)
    frames = np.random.default_rng(0).integers(0, 256, size=(n, H, W, 4), dtype=np.uint8)
    f = orc.make_uniform(W, H, blur=1, blur_radius=args.radius, edge_detect=1)
    # bound the whole run to a few minutes: cap steps by a time budget
    for _ in range(min(args.warmup, 1)):
        orc.chain_batch(frames, f, orc.CS_RAW)
    steps, t0 = 0, time.perf_counter()
    for _ in range(args.steps):
        orc.chain_batch(frames, f, orc.CS_RAW)
        steps += 1
        if time.perf_counter() - t0 > 120:
            break
    dt = time.perf_counter() - t0
    v = steps * n * PX / dt / 1e6
    sample = f"{n} frames 1920x1080 per step, {steps} steps, {nthr} host threads"
    print(json.dumps({
        "impl": "reference", "metric": "Mpixel/s, blur(r=%d)+Sobel+SSIM chain, batched 1080p RGBA8 frames" % args.radius,
        "value": v, "unit": "Mpixel/s", "n_gpus": args.gpus, "steps": steps, "warmup": min(args.warmup, 1),
        "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "batch of 1920x1080 RGBA8 frames, blur r=%d + Sobel + SSIM (BASELINE configs[3]); "
                               "CPU arm: bounded sample of %d frames per step" % (args.radius, n)},
        "cpu_baseline": {"value": v, "unit": "Mpixel/s", "cores": nthr, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is Rust+WGSL (no rustc, no wgpu adapter on this box): timed arm is the C port in oracle/",
    }))


# --------------------------------------------------------------------------------------- GPU arm

def pcie_ceilings(torch, dev, barrier, nbytes=1 << 30):
    """Pinned-memory copy rates of this rank's GPU, GB/s: H2D alone, D2H alone, both at once (each direction's rate).
    Every repetition starts behind a barrier over the ranks, so at N > 1 the figure is the rate with all GPUs of the box
    copying at once (they share the host's memory and PCIe root ports) - the ceiling the end-to-end path of N ranks
    actually has.  The FASTEST of three repetitions is reported per rank (a ceiling is the best the links did; with the
    slowest, the pipelined end-to-end path measured 8 % above its own "ceiling" at N = 8) and the caller takes the
    minimum over the ranks, since the step time is the slowest rank's."""
    h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    d_b = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def run(h2d, d2h, reps=3):
        best = float("inf")
        for rep in range(reps + 1):
            barrier()
            t0 = time.perf_counter()
            if h2d:
                with torch.cuda.stream(s1):
                    d_a.copy_(h_in, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_b, non_blocking=True)
            torch.cuda.synchronize()
            if rep:  # (the first repetition warms the path up)
                best = min(best, time.perf_counter() - t0)
        return nbytes / best / 1e9
    out = {"h2d_gbs": run(True, False), "d2h_gbs": run(False, True), "both_each_direction_gbs": run(True, True),
           "bytes": nbytes, "how": "torch pinned <-> device copies of 1 GiB, all ranks at once behind a barrier, fastest of 3 (then the minimum over ranks), wall clock around a device sync"}
    del h_in, h_out, d_a, d_b
    return out


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import numpy as np
    import torch
    import norm_b200 as nb
    from synth import make_frames

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the CUDA library is the only implementation")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_(x, op):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    def max_over_ranks(x): return reduce_(x, "MAX")

    # ---- the batch, sharded: this rank holds global frames [frame0, frame0 + n)
    total_frames = args.frames
    frame0, n = nb.plan_frames(total_frames, world, rank)
    src_t = make_frames(torch, frame0, n, H, W, dev)
    dst_t = torch.empty_like(src_t)
    ssim_t = torch.zeros(n, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    ctx = nb.Context(local_rank)
    pitch, fstride = W * 4, W * 4 * H
    src = ctx.image_wrap(src_t.data_ptr(), W, H, n, pitch, fstride)
    dst = ctx.image_wrap(dst_t.data_ptr(), W, H, n, pitch, fstride)

    def uniform(radius):
        f = nb.FeatureUniform(W, H)
        f.raw.blur, f.raw.blur_radius, f.raw.edge_detect = 1, radius, 1
        return f

    def timed(radius, steps, warmup, sample_clocks=False):
        f = uniform(radius)
        for _ in range(warmup):
            ctx.chain_async(f, src, dst, nb.NORM_CS_RAW, ssim_t.data_ptr())
        ctx.sync()
        barrier()
        sampler = ClockSampler(local_rank) if sample_clocks else None
        if sampler:
            sampler.start()
            time.sleep(0.3)
        l0, t0 = ctx.launch_count(), time.time()
        if sample_clocks:  # `ncu --profile-from-start off` then lists exactly the timed region's launches
            torch.cuda.cudart().cudaProfilerStart()
        ctx.timer_start()
        for _ in range(steps):
            ctx.chain_async(f, src, dst, nb.NORM_CS_RAW, ssim_t.data_ptr())
        ctx.timer_stop()
        ms = ctx.timer_elapsed_ms()
        ctx.sync()
        if sample_clocks:
            torch.cuda.cudart().cudaProfilerStop()
        t1 = time.time()
        barrier()
        clocks = sampler.stop(t0, t1) if sampler else None
        return max_over_ranks(ms), ctx.launch_count() - l0, clocks

    warm = max(args.warmup, 3)
    ms, launches, clocks = timed(args.radius, args.steps, warm, sample_clocks=True)
    ms_per_step = ms / args.steps
    value = total_frames * PX / (ms_per_step * 1e-3) / 1e6

    # ---- parity of the TIMED batch: 12 global frames against the CPU oracle (bytes equal, SSIM within 1e-9)
    import oracle_lib as orc
    orc.build()
    orc.lib().oracle_set_num_threads(max(1, len(os.sched_getaffinity(0)) // world))
    ss = ssim_t.cpu().numpy()
    assert np.isfinite(ss).all() and (np.abs(ss) < 1).all(), "SSIM results are not finite"
    picks = sorted({0, min(1, total_frames - 1), total_frames // 2 - 1 if total_frames > 1 else 0, total_frames - 1} |
                   set(int(x) for x in np.random.default_rng(4096).integers(0, total_frames, 8)))
    mine = [g for g in picks if frame0 <= g < frame0 + n]
    ok, max_ssim_err = 1.0, 0.0
    of = orc.make_uniform(W, H, blur=1, blur_radius=args.radius, edge_detect=1)
    for g in mine:
        a = src_t[g - frame0].cpu().numpy()[None]
        got = dst_t[g - frame0].cpu().numpy()[None]
        exp, es = orc.chain_batch(a, of, orc.CS_RAW)
        ok = min(ok, 1.0 if np.array_equal(got, exp) else 0.0)
        max_ssim_err = max(max_ssim_err, abs(float(es[0]) - float(ss[g - frame0])))
    ok = reduce_(ok, "MIN")
    max_ssim_err = reduce_(max_ssim_err, "MAX")
    checked = int(reduce_(float(len(mine)), "SUM"))
    assert ok == 1.0 and max_ssim_err <= 1e-9, f"timed batch differs from the oracle (bytes ok={ok}, SSIM err {max_ssim_err})"

    # ---- roofline of the dominant kernel, timed alone with CUDA events on its stream
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json, burst copy)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    # The step is ONE fused kernel per rank (chain_fused2_kernel<R,...>: front + blur + Sobel + SSIM moments) plus a
    # finalize launch of <10 us, so the kernel's launch duration is the CUDA-event step time measured above on
    # the library's own stream (max over ranks).  Algorithmic bytes: 4 B/px read + 4 B/px written, for
    # the pixels ONE launch processes (this rank's shard).
    n_max = int(max_over_ranks(float(n)))
    k_bytes = n_max * PX * 8.0
    kms = ms_per_step
    achieved = k_bytes / (kms * 1e-3) / 1e9
    ncu_traffic, traffic_src = None, None
    try:  # per-launch DRAM bytes from the committed ncu capture of the same kernel, scaled to this launch
        prof = json.load(open(os.path.join(ROOT, "profiles", "fused_r%d_summary.json" % args.radius)))
        ncu_traffic = prof["dram_bytes_per_pixel"] * n_max * PX
        traffic_src = "profiles/fused_r%d_summary.json (ncu --set full, dram bytes per pixel) x this launch's pixels" % args.radius
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "chain_fused2_kernel<R=%d> (front+blur+Sobel+SSIM, one pass)" % args.radius,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic,
                "traffic_source": traffic_src,
                "peak_source": peak_src, "algorithmic_bytes_per_pixel": 8, "launch_pixels": n_max * PX,
                "launch_ms": kms,
                "note": "FP32-issue-bound, not HBM-bound, at this radius: see DESIGN.md 'Roofline' and profiles/"}

    # ---- e2e: host buffers in, host buffers out, through the C ABI
    en = max(1, min(args.e2e_frames, n))
    hin, hin_p = ctx.host_alloc(en * PX * 4)
    hout, hout_p = ctx.host_alloc(en * PX * 4)
    hin_t = torch.from_numpy(hin)
    hin_t.copy_(src_t[:en].reshape(-1))  # D2H once, outside the timed region
    torch.cuda.synchronize()
    f = uniform(args.radius)
    ctx.chain_host(f, hin, hout, en, nb.NORM_CS_RAW)
    barrier()
    esteps = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(esteps):
        es = ctx.chain_host(f, hin, hout, en, nb.NORM_CS_RAW)
    e_dt = time.perf_counter() - t0
    barrier()
    e_dt = max_over_ranks(e_dt)
    en_total = int(reduce_(float(en), "SUM"))
    e2e_val = en_total * PX * esteps / e_dt / 1e6
    assert np.allclose(es, ss[:en], atol=1e-12), "host pipeline and device path disagree"
    assert np.array_equal(hout.reshape(en, H, W, 4)[0], dst_t[0].cpu().numpy()), "host pipeline output differs"
    ctx.host_free(hin_p), ctx.host_free(hout_p)
    pcie = pcie_ceilings(torch, dev, barrier)
    for k in ("h2d_gbs", "d2h_gbs", "both_each_direction_gbs"):  # the slowest rank's figure
        pcie[k] = reduce_(pcie[k], "MIN")
    barrier()
    # the chain through host buffers moves 4 B/px each way at once: its ceiling is the both-directions rate
    pcie["e2e_gbs_each_direction"] = e2e_val / world * 4.0 / 1e3
    pcie["e2e_fraction_of_ceiling"] = pcie["e2e_gbs_each_direction"] / pcie["both_each_direction_gbs"]

    sweep = None
    if not args.no_sweep:
        sweep = {}
        for r in (3, 7, 39):
            k = max(3, args.steps // 2)
            sms, _, _ = timed(r, k, 2)
            sweep[str(r)] = total_frames * PX / (sms / k * 1e-3) / 1e6

    # ---- BASELINE configs[4]: one giant image, row-band sharded (needs peers: N > 1)
    band = None
    src.destroy(), dst.destroy()
    del src_t, dst_t
    torch.cuda.empty_cache()
    if world > 1 and not args.no_band:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from band_run import band_benchmark
        band = band_benchmark(torch, dist, ctx, local_rank, rank, world, args.band_size, args.band_size, args.radius,
                              steps=20, warmup=3, verify="full")  # 20 steps queued back to back, one sync at the end
        band["config"] = "BASELINE.json configs[4]"

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_chain_rate(args.radius)

    if rank == 0:
        info = ctx.device_info()
        line = {
            "metric": "Mpixel/s, blur(r=%d)+Sobel+SSIM chain, batched 1080p RGBA8 frames" % args.radius,
            "value": value, "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "ONE batch of %d 1920x1080 RGBA8 frames (BASELINE configs[3]) sharded over %d GPU(s) with "
                                   "norm_plan_frames, Gaussian blur r=%d -> Sobel -> SSIM vs source, colour-space raw"
                                   % (total_frames, world, args.radius),
                       "frames_total": total_frames, "frames_per_gpu_max": n_max, "radius": args.radius,
                       "l2": "a rank's shard (%.1f GB) exceeds the 126 MB L2; no flush needed" % (n_max * PX * 4 / 1e9),
                       "device": info["name"], "sm_count": info["sm_count"]},
            "roofline": roofline,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_val, "unit": "Mpixel/s", "h2d_bytes_per_step": en * PX * 4,
                    "d2h_bytes_per_step": en * PX * 4 + en * 8, "frames_per_gpu": en, "steps": esteps,
                    "api": "norm_cuda_chain_host (pinned host RGBA8 in, RGBA8 + SSIM out)", "pcie": pcie},
            "gpu_launches": launches,
            "clocks": clocks,
            "parity_checked_frames": checked,
            "parity": {"frames": picks, "bytes_equal": True, "max_ssim_abs_err": max_ssim_err,
                       "against": "oracle/libnorm_oracle.so on the frames of the timed batch"},
        }
        if sweep:
            line["radius_sweep_mpixel_s"] = sweep
        if band:
            line["band"] = band
        print(json.dumps(line))
    ctx.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
