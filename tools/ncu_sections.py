#!/usr/bin/env python
"""Summarise an ncu report of a kernel: headline metrics, stall mix, and per-section (between
barrier instructions) instruction / stall shares.  usage: tools/ncu_sections.py <rep> [pixels]"""
import csv
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
px = float(sys.argv[2]) if len(sys.argv) > 2 else 128 * 1920 * 1080
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
d = dict(zip(rows[0], rows[2]))
for k in ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum',
          'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
          'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
          'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
          'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
          'smsp__warps_eligible.avg.per_cycle_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
          'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum']:
    print(f"{k:80s} {d.get(k)}")
st = [(float(v.replace(',', '')), h) for h, v in d.items() if 'average_warps_issue_stalled' in h and 'per_issue_active' in h]
st.sort(reverse=True)
for v, h in st[:7]:
    print(f'{v:8.3f} {h}')
print('lane-instr/px', float(d['smsp__inst_executed.sum'].replace(',', '')) * 32 / px)
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
names = ['Instructions Executed', '# Samples', 'stall_barrier', 'stall_long_sb', 'stall_math', 'stall_wait', 'stall_short_sb',
         'stall_not_selected', 'stall_selected', 'stall_no_inst', 'stall_dispatch', 'stall_branch_resolving', 'stall_mio']
cols = {h: hdr.index(h) for h in ['Source'] + names}
inst = [r for r in rows[hi + 1:] if len(r) > cols['# Samples']]
tot = defaultdict(float)
acc = defaultdict(float)
segs = []
for r in inst:
    src = r[cols['Source']]
    for k in names:
        v = float(r[cols[k]] or 0)
        acc[k] += v
        tot[k] += v
    if 'BAR' in src or 'EXIT' in src or 'SETMAXREG' in src.upper():
        segs.append((src.strip()[:44], dict(acc)))
        acc = defaultdict(float)
ts, ti = tot['# Samples'], tot['Instructions Executed']
for name, a in segs:
    if a['# Samples'] / ts > 0.004:
        print(f"{a['# Samples']/ts*100:5.1f}% samp {a['Instructions Executed']/ti*100:5.1f}% inst | bar {a['stall_barrier']/ts*100:4.1f} "
              f"lsb {a['stall_long_sb']/ts*100:4.1f} wait {a['stall_wait']/ts*100:4.1f} math {a['stall_math']/ts*100:4.1f} "
              f"ssb {a['stall_short_sb']/ts*100:4.1f} notsel {a['stall_not_selected']/ts*100:4.1f} sel {a['stall_selected']/ts*100:4.1f} "
              f"noinst {a['stall_no_inst']/ts*100:4.1f} disp {a['stall_dispatch']/ts*100:4.1f} -> {name}")
