#!/bin/bash
# Builds development variants of the library for A/B runs on the GPU box:
#   tools/ab/build_variants.sh name1 "flags1" name2 "flags2" ...   ->  tools/ab/<name>.so
# (BASELINE radii only; no --split-compile: it changes register allocation)
set -e
cd "$(dirname "$0")/../.."
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  NORM_ALLOW_SPILLS=1 NORM_NVCC_EXTRA="-DNORM_NO_EXTRA_RADII $flags" python build.py > /dev/null
  cp norm_b200/libnorm_cuda.so tools/ab/$name.so
  grep -A3 "chain_fused2_kernelILi21ELb0ELb0ELi0" norm_b200/build.log | grep -E "spill" | sed "s/^/$name: /"
done
