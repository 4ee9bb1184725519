#!/usr/bin/env python
"""Stall samples of an ncu report's kernel by source region: ncu_roles.py file.ncu-rep [top-n]
Prints the top stalled SASS instructions and an opcode histogram (executions per warp)."""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = rows[1]; data = rows[2:]
iS = hdr.index("# Samples"); iSrc = hdr.index("Source"); iE = hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[iS] or 0) for r in data)
print("kernel:", rows[0][1][:100]); print("total samples", tot)
for idx, r in sorted(enumerate(data), key=lambda t: -int(t[1][iS] or 0))[:topn]:
    st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f"{idx:5d} {100*int(r[iS])/tot:5.1f}% ex={r[iE]:>10s} {r[iSrc][:64]:64s} {st}")
d = collections.Counter()
for i in stall_cols:
    d[hdr[i][6:]] = sum(int(r[i] or 0) for r in data)
print("stall mix:", [(k, round(100 * v / tot, 1)) for k, v in d.most_common(8)])
