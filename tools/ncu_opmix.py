#!/usr/bin/env python
"""Opcode mix (lane-instructions per pixel) of every distinct kernel of an ncu report captured with --set full.
usage: tools/ncu_opmix.py <report.ncu-rep> <pixels per launch> [kernel-name substring]"""
import collections
import csv
import subprocess
import sys

rep, px = sys.argv[1], float(sys.argv[2])
want = sys.argv[3] if len(sys.argv) > 3 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
i, seen = 0, set()
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]
        hdr = rows[i + 1]
        col = {h: k for k, h in enumerate(hdr)}
        j = i + 2
        ops, tot = collections.Counter(), 0.0
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            r = rows[j]
            if len(r) > col["Instructions Executed"] and r[col["Source"]].strip():
                s = r[col["Source"]].strip()
                op = (s.split()[1] if s.startswith("@") else s.split()[0]).split(".")[0]
                n = float(r[col["Instructions Executed"]] or 0)
                ops[op] += n
                tot += n
            j += 1
        if name not in seen and want in name:
            seen.add(name)
            print(name)
            print(f"  {tot * 32 / px:.1f} lane-instr/px: " + ", ".join(f"{k} {v * 32 / px:.1f}" for k, v in ops.most_common(28)))
        i = j
    else:
        i += 1
