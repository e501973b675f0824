#!/usr/bin/env python
"""One line block per distinct kernel of an ncu report (--set full): duration, DRAM bytes, issue slots, occupancy,
pipes and the top stall reasons.  usage: tools/ncu_stage_summary.py <report.ncu-rep> [pixels per launch]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
px = float(sys.argv[2]) if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
seen = set()


def num(d, k):
    try:
        return float(d.get(k, "nan").replace(",", ""))
    except ValueError:
        return float("nan")


for r in rows[2:]:
    d = dict(zip(hdr, r))
    name = d.get("Kernel Name", "?")
    if name in seen:
        continue
    seen.add(name)
    u = dict(zip(hdr, units))
    dur = num(d, "gpu__time_duration.sum")
    dur_us = dur / 1e3 if u.get("gpu__time_duration.sum") == "ns" else dur * (1e3 if u.get("gpu__time_duration.sum") == "ms" else 1)
    rd, wr = num(d, "dram__bytes_read.sum"), num(d, "dram__bytes_write.sum")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rd *= scale.get(u.get("dram__bytes_read.sum"), 1)
    wr *= scale.get(u.get("dram__bytes_write.sum"), 1)
    inst = num(d, "smsp__inst_executed.sum")
    print(name)
    print(f"  duration {dur_us:9.1f} us   DRAM read {rd/1e6:9.1f} MB  write {wr/1e6:9.1f} MB  -> {(rd+wr)/dur_us/1e3:7.0f} GB/s "
          f"({num(d, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.1f} % of ncu's DRAM peak)")
    line = (f"  issue slots busy {num(d, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):5.1f} %   warps active "
            f"{num(d, 'sm__warps_active.avg.pct_of_peak_sustained_active'):5.1f} %   eligible warps/cycle "
            f"{num(d, 'smsp__warps_eligible.avg.per_cycle_active'):.2f}   registers {d.get('launch__registers_per_thread')}   "
            f"warp-instructions {inst:.3e}")
    if px:
        line += f" = {inst * 32 / px:.1f} lane-instr/px, DRAM {(rd + wr) / px:.2f} B/px"
    print(line)
    print(f"  pipes: fma {num(d, 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active'):.1f} %  alu "
          f"{num(d, 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'):.1f} %  lsu "
          f"{num(d, 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active'):.1f} %  xu "
          f"{num(d, 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active'):.1f} %")
    st = [(num(d, h), h) for h in hdr if "average_warps_issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    st = sorted([x for x in st if x[0] == x[0]], reverse=True)[:4]
    print("  stalls per issue: " + ", ".join(f"{h.split('issue_stalled_')[1].split('_per_issue')[0]} {v:.2f}" for v, h in st))
