#!/usr/bin/env python
"""top SASS instructions by stall samples from an ncu report: python tools/ncu_top.py <rep> [n] [kernel-index]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = txt.split('"Kernel Name",')
which = int(sys.argv[3]) if len(sys.argv) > 3 else 1
blk = blocks[which]
lines = blk.split("\n")
print("kernel:", lines[0][:120])
r = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
h = r[0]; ci = {k: i for i, k in enumerate(h)}
rows = [x for x in r[1:] if len(x) == len(h)]
tot = sum(float(x[ci["# Samples"]] or 0) for x in rows)
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
print("total samples", tot)
agg = {k: sum(float(x[ci[k]] or 0) for x in rows) for k in stalls}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
order = sorted(range(len(rows)), key=lambda i: -float(rows[i][ci["# Samples"]] or 0))
for i in order[:n]:
    x = rows[i]
    top = sorted(((float(x[ci[k]] or 0), k) for k in stalls), reverse=True)[:2]
    print(f'{x[ci["# Samples"]]:>7} exec={x[ci["Instructions Executed"]]:>9} #{i:<5} {x[ci["Source"]].strip()[:90]:<90} {top}')
