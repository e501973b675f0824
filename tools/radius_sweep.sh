#!/bin/bash
# chain and blur-only throughput for every radius the reference UI reaches (3..39 step 2), 128 1080p frames
for r in $(seq 3 2 39); do
  python tools/prof_chain.py --frames 128 --reps 3 --radius $r
  python tools/prof_chain.py --frames 128 --reps 3 --radius $r --gray 1
  python tools/prof_chain.py --frames 128 --reps 3 --radius $r --blur-only
done
