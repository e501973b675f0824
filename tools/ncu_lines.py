#!/usr/bin/env python
"""Attribute ncu per-instruction counters to CUDA source lines.

usage: tools/ncu_lines.py <report.ncu-rep> <mangled-kernel-substring> [top]
Joins `ncu --page source --csv` (SASS view: executed instructions + stall samples per
instruction, in program order) with `nvdisasm -g` line markers of the same kernel in
norm_b200/libnorm_cuda.so (instruction order is identical).
"""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40

raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
metric = os.environ.get("NCU_METRIC", "Instructions Executed")
ci, cs, csrc = hdr.index(metric), hdr.index("# Samples"), hdr.index("Source")
inst = [(r[csrc].strip(), float(r[ci] or 0), float(r[cs] or 0)) for r in rows[hi + 1:] if len(r) > ci]

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "norm_b200", "libnorm_cuda.so")], cwd=tmp,
               capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
lines, cur, on = [], None, False
for ln in sass.splitlines():
    if ln.startswith("\t.section\t.text."):
        on = kern in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
    if m:
        lines.append((cur, m.group(1).strip()))
if len(lines) != len(inst):
    print(f"warning: {len(lines)} disassembled vs {len(inst)} profiled instructions", file=sys.stderr)
agg = defaultdict(lambda: [0.0, 0.0, defaultdict(float)])
for (loc, txt), (_, n, s) in zip(lines, inst):
    a = agg[loc]
    a[0] += n
    a[1] += s
    op = txt.split()[0] if not txt.startswith("@") else txt.split()[1]
    a[2][op.split(".")[0]] += n
tot = sum(a[0] for a in agg.values()) or 1
tots = sum(a[1] for a in agg.values()) or 1
src_cache = {}
print(f"total warp-instructions {tot:.3e}, samples {tots:.0f}")
for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    text = ""
    if loc:
        for d in ("norm_b200/csrc",):
            p = os.path.join(ROOT, d, loc[0])
            if os.path.exists(p):
                src_cache.setdefault(p, open(p).read().splitlines())
                text = src_cache[p][loc[1] - 1].strip()[:70]
    ops = " ".join(f"{k}:{v / tot * 100:.1f}" for k, v in sorted(a[2].items(), key=lambda kv: -kv[1])[:4])
    print(f"{a[0] / tot * 100:5.1f}% inst {a[1] / tots * 100:5.1f}% stall  {loc[0] if loc else '?'}:{loc[1] if loc else 0:<4} {text:70s} | {ops}")
