#!/bin/bash
# tools/ab/libs.sh "lib1 lib2 ..." "args1" "args2" ...: prof_chain for every (args, lib) pair
libs=$1; shift
for extra in "$@"; do
  for l in $libs; do
    python tools/prof_chain.py --frames 256 --reps 5 $extra --lib tools/ab/$l.so | sed "s/^/$l /"
  done
done
