#!/bin/bash
# A/B of the blur-only (MODE 1) kernels: tools/ab/base.so vs the in-tree build
for extra in "--radius 21" "--radius 19" "--radius 13" "--radius 7" "--radius 21 --gray 1"; do
  python tools/prof_chain.py --frames 256 --reps 5 --blur-only $extra --lib tools/ab/base.so | sed 's/^/base /'
  python tools/prof_chain.py --frames 256 --reps 5 --blur-only $extra | sed 's/^/new  /'
done
