#!/bin/bash
# A/B of the chain (MODE 0) kernels: tools/ab/base.so vs the in-tree build
for extra in "--radius 21" "--radius 21" "--radius 7" "--radius 3" "--radius 21 --gray 1"; do
  python tools/prof_chain.py --frames 256 --reps 5 $extra --lib tools/ab/base.so | sed 's/^/base /'
  python tools/prof_chain.py --frames 256 --reps 5 $extra | sed 's/^/new  /'
done
