#!/usr/bin/env python
"""Headline benchmark: Mpixel/s of the blur + Sobel + SSIM chain on batched 1080p RGBA8 frames.

    python bench.py [--gpus N] [--steps K] [--warmup W]                       (this repo's CUDA path)
    python bench.py --impl reference [--gpus N] [--steps K] [--warmup W]      (CPU arm)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[3]): frames of 1920x1080 RGBA8, Gaussian blur at the reference's
default radius 21 -> Sobel line detection -> global SSIM of the result against the source frame.
Frames are independent, so they shard across GPUs with no data-path collective: every rank
holds `--frames` frames (weak scaling; configs[3]'s 4096 frames is the per-GPU default) and
`value` = all ranks' pixels / max-over-ranks device time.

A step = one pass of the chain over the rank's whole batch, inputs resident in HBM.  The batch
(34 GB) is ~270x the 126 MB L2, so no explicit L2 flush is needed between steps.
`e2e` = the same chain through norm_cuda_chain_host on pinned HOST buffers (H2D + D2H in the
timed region) over a bounded subset of the batch.
The CPU arm / cpu_baseline time oracle/libnorm_oracle.so (a C restatement of the reference; the
reference itself is Rust + WGSL and cannot be built or run here) on all host cores.
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
    ap.add_argument("--frames", type=int, default=4096, help="frames per GPU")
    ap.add_argument("--radius", type=int, default=21)
    ap.add_argument("--e2e-frames", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep", action="store_true", help="also report radius 3/7/39 (extra keys)")
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

def cpu_chain_rate(radius, seconds=12.0):
    """Oracle (C port of the reference path) on all host threads, bounded sample. -> dict"""
    import numpy as np
    import oracle_lib as orc
    orc.build()
    nthr = orc.lib().oracle_num_threads()
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
    nthr = orc.lib().oracle_num_threads()
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
        "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "batch of 1920x1080 RGBA8 frames, blur r=%d + Sobel + SSIM (BASELINE configs[3]); "
                               "CPU arm: bounded sample of %d frames per step" % (args.radius, n)},
        "cpu_baseline": {"value": v, "unit": "Mpixel/s", "cores": nthr, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is Rust+WGSL (no rustc, no wgpu adapter on this box): timed arm is the C port in oracle/",
    }))


# --------------------------------------------------------------------------------------- GPU arm

def make_frames(torch, n, device, seed):
    """n frames (n,H,W,4) uint8 on device: PRNG noise blended with a per-frame smooth gradient."""
    out = torch.empty((n, H, W, 4), dtype=torch.uint8, device=device)
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    ys = torch.linspace(0, 1, H, device=device).view(1, H, 1, 1)
    xs = torch.linspace(0, 1, W, device=device).view(1, 1, W, 1)
    step = 32
    for i in range(0, n, step):
        m = min(step, n - i)
        noise = torch.randint(0, 256, (m, H, W, 4), dtype=torch.uint8, device=device, generator=g)
        ph = (torch.arange(i, i + m, device=device, dtype=torch.float32) * 0.37).view(m, 1, 1, 1)
        ch = torch.arange(4, device=device, dtype=torch.float32).view(1, 1, 1, 4)
        grad = 127.5 + 127.5 * torch.sin(6.0 * xs * (1 + 0.1 * ch) + 4.0 * ys + ph)
        out[i:i + m] = ((noise.to(torch.float32) + grad) * 0.5).to(torch.uint8)
        out[i:i + m, :, :, 3] = 255
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

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = args.frames
    src_t = make_frames(torch, n, dev, seed=0x6E6F726D + rank)
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
        ctx.timer_start()
        for _ in range(steps):
            ctx.chain_async(f, src, dst, nb.NORM_CS_RAW, ssim_t.data_ptr())
        ctx.timer_stop()
        ms = ctx.timer_elapsed_ms()
        ctx.sync()
        t1 = time.time()
        barrier()
        clocks = sampler.stop(t0, t1) if sampler else None
        return max_over_ranks(ms), ctx.launch_count() - l0, clocks

    ms, launches, clocks = timed(args.radius, args.steps, max(args.warmup, 3), sample_clocks=True)
    ms_per_step = ms / args.steps
    value = world * n * PX / (ms_per_step * 1e-3) / 1e6

    # sanity: the timed work really happened and is sane (SSIM of an edge map vs its source is small but finite)
    ss = ssim_t.cpu().numpy()
    assert np.isfinite(ss).all() and (np.abs(ss) < 1).all(), "SSIM results are not finite"

    # ---- roofline of the dominant kernel, timed alone with CUDA events on its stream
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json, burst copy)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    # The step is ONE fused kernel (chain_fused2_kernel<R,...>: front + blur + Sobel + SSIM moments) plus a
    # finalize launch of <10 us, so the kernel's launch duration is the CUDA-event step time measured above on
    # the library's own stream.  Algorithmic bytes: 4 B/px source read + 4 B/px result written.
    k_bytes = n * PX * 8.0
    kms = ms_per_step  # max over ranks
    achieved = k_bytes / (kms * 1e-3) / 1e9
    ncu_traffic = None
    try:  # per-launch DRAM bytes from the committed ncu capture of the same kernel, scaled to this launch
        prof = json.load(open(os.path.join(ROOT, "profiles", "fused_r%d_summary.json" % args.radius)))
        ncu_traffic = prof["dram_bytes_per_pixel"] * n * PX
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "chain_fused2_kernel<R=%d> (front+blur+Sobel+SSIM, one pass)" % args.radius,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic,
                "peak_source": peak_src, "algorithmic_bytes_per_pixel": 8, "launch_pixels": n * PX,
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
    e2e_val = world * en * PX * esteps / e_dt / 1e6
    assert np.allclose(es, ss[:en], atol=1e-12), "host pipeline and device path disagree"
    ctx.host_free(hin_p), ctx.host_free(hout_p)

    sweep = None
    if args.sweep:
        sweep = {}
        for r in (3, 7, 39):
            sms, _, _ = timed(r, max(3, args.steps // 2), 2)
            sweep[str(r)] = world * n * PX / (sms / max(3, args.steps // 2) * 1e-3) / 1e6

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_chain_rate(args.radius)

    if rank == 0:
        info = ctx.device_info()
        line = {
            "metric": "Mpixel/s, blur(r=%d)+Sobel+SSIM chain, batched 1080p RGBA8 frames" % args.radius,
            "value": value, "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "batch of %d 1920x1080 RGBA8 frames per GPU (BASELINE configs[3]), Gaussian blur "
                                   "r=%d -> Sobel -> SSIM vs source, frame-sharded, colour-space raw" % (n, args.radius),
                       "frames_per_gpu": n, "radius": args.radius,
                       "l2": "inputs (%.1f GB per GPU) exceed the 126 MB L2; no flush needed" % (n * PX * 4 / 1e9),
                       "device": info["name"], "sm_count": info["sm_count"]},
            "roofline": roofline,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_val, "unit": "Mpixel/s", "h2d_bytes_per_step": en * PX * 4,
                    "d2h_bytes_per_step": en * PX * 4 + en * 8, "frames_per_gpu": en, "steps": esteps,
                    "api": "norm_cuda_chain_host (pinned host RGBA8 in, RGBA8 + SSIM out)"},
            "gpu_launches": launches,
            "clocks": clocks,
        }
        if sweep:
            line["radius_sweep_mpixel_s"] = sweep
        print(json.dumps(line))
    src.destroy(), dst.destroy()
    ctx.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
