#!/bin/bash
# A/B of several builds of the library: tools/ab/variants.sh name1 name2 ...   (tools/ab/<name>.so)
# Prints Gpixel/s of the chain for r = 21 / 7 / 3 on RGB and r = 21 after grayscale, two rounds each.
for rep in 1 2; do
  for n in "$@"; do
    for extra in "--radius 21" "--radius 7" "--radius 3" "--radius 21 --gray 1"; do
      python tools/prof_chain.py --frames 256 --reps 5 $extra --lib tools/ab/$n.so | sed "s/^/$n  /"
    done
  done
done
