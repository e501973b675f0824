#!/usr/bin/env python
"""Builds the in-tree native artefacts:
  norm_b200/libnorm_cuda.so   nvcc, sm_100a only (the product)
  oracle/libnorm_oracle.so    gcc (the CPU checker; test infrastructure)
Both .so files are git-ignored and travel to the GPU box with the gpurun snapshot.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "norm_b200", "csrc")
LIB = os.path.join(ROOT, "norm_b200", "libnorm_cuda.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-fmad=false",                      # no silent FMA contraction: every fma in the kernels is explicit
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-Wall",
    "-Xptxas", "-v",
    "-shared", "-cudart", "static",
]
# Development builds: NORM_NVCC_EXTRA="--split-compile 0 -DNORM_NO_EXTRA_RADII" compiles in about 30 s instead of
# 3.4 min.  Not the default: --split-compile changes register allocation in a few kernels (measured: grayscale
# chain r=21 158 -> 148 Gpixel/s, RGBA blur r=21 80 -> 75); results stay bit-exact either way.


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_cuda(force=False, verbose=False):
    srcs = [os.path.join(CSRC, n) for n in sorted(os.listdir(CSRC)) if n != "norm_nccl.cu"]
    srcs += [os.path.join(ROOT, "include", "norm_cuda.h"), os.path.abspath(__file__)]
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    extra = os.environ.get("NORM_NVCC_EXTRA", "").split()
    cmd = [nvcc, "-ccbin", "/usr/bin/g++"] + NVCC_FLAGS + extra + ["-o", LIB, os.path.join(CSRC, "norm_cuda.cu")]
    # a library built with other flags (e.g. the quick development build) counts as stale
    flags_file = LIB + ".flags"
    same_flags = os.path.exists(flags_file) and open(flags_file).read() == " ".join(cmd)
    if not (force or not same_flags or _stale(LIB, srcs)):
        return LIB
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode == 0:
        with open(flags_file, "w") as f:
            f.write(" ".join(cmd))
    log = os.path.join(ROOT, "norm_b200", "build.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed, see " + log)
    if not os.environ.get("NORM_ALLOW_SPILLS"):  # (experiments with other register budgets set it)
        check_spills(r.stdout + r.stderr)
    return LIB


SPILL_BYTES_ALLOWED = 24


def check_spills(ptxas_log):
    """The pipelined chain kernel (chain_fused2_kernel<R, GRAY, FRONT, MODE 0>: the headline path) must not spill:
    its register budgets are tuned per warp group and a spill in a hot loop costs tens of percent.  ptxas -v reports
    spills per entry function; more than SPILL_BYTES_ALLOWED bytes of spill stores in a MODE 0 instantiation fails the
    build.  (Up to 24 bytes are tolerated: at some radii ptxas parks a few loop-invariant per-thread values - staging
    column offsets computed once before the item loop - on the stack and reloads them once per batch or H item; seen
    in the SASS of r = 29, group A, 20 bytes.  A spilled accumulator or tap shows up as hundreds of bytes.)"""
    import re
    bad, cur = [], None
    for line in ptxas_log.splitlines():
        m = re.search(r"Function properties for (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"(\d+) bytes spill stores", line)
        if m and cur and "chain_fused2_kernel" in cur and cur.endswith("ELi0EEEvNS_11FusedParamsE") and int(m.group(1)) > SPILL_BYTES_ALLOWED:
            bad.append((cur, int(m.group(1))))
    if bad:
        raise RuntimeError("register spills in the pipelined chain kernel: " + ", ".join(f"{n} ({b} B)" for n, b in bad))


NCCL_LIB = os.path.join(ROOT, "norm_b200", "libnorm_nccl.so")


def build_nccl(force=False):
    """libnorm_nccl.so: the SSIM-moments all-reduce, linked against the system NCCL (libnccl.so.2;
    inside a torch process the already-loaded bundled NCCL with the same SONAME is reused)."""
    src = os.path.join(CSRC, "norm_nccl.cu")
    deps = [src, os.path.join(ROOT, "include", "norm_nccl.h"), os.path.abspath(__file__)]
    if not (force or _stale(NCCL_LIB, deps)):
        return NCCL_LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    env = dict(os.environ)
    env.pop("CC", None), env.pop("CXX", None)
    cmd = [nvcc, "-ccbin", "/usr/bin/g++", "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "-o", NCCL_LIB, src, "-lnccl"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed for norm_nccl.cu")
    return NCCL_LIB


def build_oracle(force=False):
    args = ["make", "-C", os.path.join(ROOT, "oracle"), "-s"]
    if force:
        args.append("-B")
    subprocess.check_call(args)
    return os.path.join(ROOT, "oracle", "libnorm_oracle.so")


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_cuda(force=force, verbose="-v" in sys.argv))
    print(build_nccl(force=force))
    print(build_oracle(force=force))
