#!/usr/bin/env python
"""Print the metrics that matter from an ncu report: ncu_summary.py file.ncu-rep [kernel-regex]"""
import csv, subprocess, sys, io, re
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
want = re.compile(r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum$|gpu__dram_throughput.avg.pct|launch__registers_per_thread$|launch__grid_size|launch__block_size|launch__occupancy_limit|sm__warps_active.avg.pct|sm__inst_executed_pipe_(alu|fma|fmaheavy|fmalite|lsu|fp64|tensor.*|uniform|xu|cbu|adu).sum.pct|sm__pipe_(alu|fma|fmaheavy|fp64|tensor.*)_cycles_active.avg.pct_of_peak_sustained_active|smsp__issue_active.avg.pct|smsp__inst_executed.sum$|smsp__average_warps_issue_stalled_.*_per_issue_active|sm__cycles_elapsed.max$|lts__t_sectors_srcunit_tex_op_read.sum$|lts__t_bytes.sum$|l1tex__t_bytes.sum$|sm__throughput.avg.pct|lts__throughput.avg.pct|l1tex__throughput.avg.pct|smsp__cycles_active.avg$|sm__inst_executed_pipe_tensor.*|sm__pipe_tensor.*)")
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    if len(sys.argv) > 2 and not re.search(sys.argv[2], name):
        continue
    print("==", name[:110], "id", r[0])
    for h, u, v in zip(hdr, units, r):
        if want.search(h):
            try:
                fv = float(v.replace(",", ""))
            except ValueError:
                continue
            if "stalled" in h and fv < 0.05:
                continue
            print(f"   {h:95s} {u:12s} {v}")
