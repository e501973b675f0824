#!/usr/bin/env python
"""Per-warp-group view of an ncu capture of chain_fused2_kernel: splits the SASS of the source page at the
group boundaries (the setmaxnreg instructions), and prints per group the executed warp instructions,
the opcode mix in lane-instructions per pixel and the stall-sample breakdown.

usage: tools/ncu_groups.py <report.ncu-rep> <pixels per launch>
"""
import collections
import csv
import subprocess
import sys

rep, px = sys.argv[1], float(sys.argv[2])
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) > col["stall_wait"]]


def f(r, k):
    return float(r[col[k]] or 0)


cuts = [i for i, r in enumerate(data) if "USETMAXREG" in r[col["Source"]] or "SETMAXREG" in r[col["Source"]]]
print("setmaxnreg at", cuts, "of", len(data))
bounds = [0] + cuts + [len(data)]
stalls = [h for h in hdr if h.startswith("stall_") and "Not" not in h]
for gi in range(len(bounds) - 1):
    a, b = bounds[gi], bounds[gi + 1]
    ops, tot, st = collections.Counter(), 0.0, collections.Counter()
    for r in data[a:b]:
        s = r[col["Source"]].strip()
        op = (s.split()[1] if s.startswith("@") else s.split()[0]).split(".")[0]
        n = f(r, "Instructions Executed")
        ops[op] += n
        tot += n
        for k in stalls:
            st[k] += f(r, k)
    if tot == 0:
        continue
    print(f"\nsegment {gi} [{a},{b}): {tot:.3e} warp-instr = {tot * 32 / px:.1f} lane-instr/px, samples {sum(st.values()):.0f}")
    print("  ops/px:", ", ".join(f"{k} {v * 32 / px:.1f}" for k, v in ops.most_common(24)))
    print("  stalls:", ", ".join(f"{k[6:]} {v:.0f}" for k, v in st.most_common(8)))
