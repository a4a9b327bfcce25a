#!/usr/bin/env python3
"""Short read of an .ncu-rep: duration, DRAM / issue / pipe utilisation, stall mix, per-opcode instruction counts.
usage: ncu_brief.py report.ncu-rep [launch-index]"""
import csv
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
li = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
h, r = raw[0], raw[2 + li]
m = dict(zip(h, r))
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
for k in keys:
    if k in m:
        print(f"{k} = {m[k]}")
st = {k.split("issue_stalled_")[1].split("_per_")[0]: float(v) for k, v in m.items() if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("per_issue_active.ratio")}
print("stalls/issue:", ", ".join(f"{k} {v:.2f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:8]))
src = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout.splitlines()))
hi = next(i for i, row in enumerate(src) if "Source" in row and "# Instructions Executed" in " ".join(row) or "Instructions Executed" in row)
hh = src[hi]
cs, ce = hh.index("Source"), hh.index("Instructions Executed")
ops, tot = Counter(), 0
for row in src[hi + 1:]:
    if len(row) <= ce or not row[ce].isdigit():
        continue
    t = row[cs].split()
    op = (t[1] if t and t[0].startswith("@") else t[0]).split(".")[0] if t else "?"
    ops[op] += int(row[ce])
    tot += int(row[ce])
print("warp instructions", tot, "static", len(src) - hi - 1)
print(", ".join(f"{o} {100 * c / tot:.1f}%" for o, c in ops.most_common(16)))
