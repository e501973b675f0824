#!/bin/bash
# A/B of two builds of the library on the chain workload: tools/ab/base.so vs the in-tree build
for r in 21 7 3; do
  for rep in 1 2; do
    python tools/prof_chain.py --frames 256 --reps 5 --radius $r --lib tools/ab/base.so | sed 's/^/base /'
    python tools/prof_chain.py --frames 256 --reps 5 --radius $r | sed 's/^/new  /'
  done
done
python tools/prof_chain.py --frames 256 --reps 5 --gray 1 --lib tools/ab/base.so | sed 's/^/base /'
python tools/prof_chain.py --frames 256 --reps 5 --gray 1 | sed 's/^/new  /'
for extra in "--radius 21" "--radius 7" "--radius 21 --gray 1"; do
  python tools/prof_chain.py --frames 256 --reps 5 --blur-only $extra --lib tools/ab/base.so | sed 's/^/base /'
  python tools/prof_chain.py --frames 256 --reps 5 --blur-only $extra | sed 's/^/new  /'
done
