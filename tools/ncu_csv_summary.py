#!/usr/bin/env python
"""Summary of an ncu capture exported on the GPU box as CSV (ncu -i x.ncu-rep --page raw --csv / --page source --csv):
launch facts, pipe utilisation, stall reasons per issue, and executed instructions / stall samples per opcode.
    ncu_csv_summary.py raw.csv [source.csv] [units]      (units: divide the instruction counts by this many work units)"""
import collections
import csv
import re
import sys

raw = list(csv.reader(open(sys.argv[1])))
hdr, units_row, vals = raw[0], raw[1], raw[2]
m = dict(zip(hdr, vals))
keys = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.sum.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.max"]
u = dict(zip(hdr, units_row))
for k in keys:
    if k in m:
        print(f"{k:70s} {u.get(k, ''):14s} {m[k]}")
st = []
for h, v in zip(hdr, vals):
    if "average_warps_issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h:
        st.append((float(v), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
print("stall cycles per issue:", ", ".join(f"{h} {v:.2f}" for v, h in sorted(st, reverse=True)[:9]))
if len(sys.argv) > 2:
    div = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    by, samp, tot = collections.Counter(), collections.Counter(), 0
    for r in list(csv.reader(open(sys.argv[2])))[2:]:
        try:
            ex, sm = int(r[5]), int(r[2])
        except (ValueError, IndexError):
            continue
        op = re.sub(r"^@!?U?P\d+\s+", "", r[1].strip()).split()[0].split(".")[0]
        by[op] += ex
        samp[op] += sm
        tot += ex
    ts = max(1, sum(samp.values()))
    print(f"warp instructions executed: {tot / div:.1f}" + (" per unit" if div != 1.0 else ""))
    for op, c in by.most_common(16):
        print(f"   {op:10s} {c / div:10.1f}   stall samples {100 * samp[op] / ts:5.1f} %")
