#!/usr/bin/env python
"""profiles/fused_r<R>_summary.json from an `ncu --set full` capture of the fused chain kernel
(tools/prof_chain.py --frames 64 --radius R).  usage: tools/ncu_fused_summary.py <report.ncu-rep> <R> <pixels> <out.json>"""
import csv
import json
import subprocess
import sys

rep, radius, px, out = sys.argv[1], int(sys.argv[2]), float(sys.argv[3]), sys.argv[4]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
d, u = dict(zip(hdr, vals)), dict(zip(hdr, units))
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
f = lambda k: float(d[k].replace(",", ""))
rd = f("dram__bytes_read.sum") * scale.get(u["dram__bytes_read.sum"], 1)
wr = f("dram__bytes_write.sum") * scale.get(u["dram__bytes_write.sum"], 1)
keys = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__icc_request_hit_rate.pct"]
json.dump({"kernel": d["Kernel Name"], "capture": "ncu --set full --clock-control none --import-source on; tools/prof_chain.py --frames 64 --radius %d" % radius,
           "pixels": px, "dram_read_bytes": rd, "dram_write_bytes": wr, "dram_bytes_per_pixel": (rd + wr) / px,
           "lane_instructions_per_pixel": f("smsp__inst_executed.sum") * 32 / px,
           "metrics": {k: d.get(k, "") + " " + u.get(k, "") for k in keys}}, open(out, "w"), indent=1)
print(open(out).read())
