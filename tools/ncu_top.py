#!/usr/bin/env python
"""top SASS instructions by stall samples from an ncu report: python tools/ncu_top.py <rep> [n] [kernel-name-regex]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cmd = ["ncu", "-i", rep, "--page", "source", "--csv"]
if len(sys.argv) > 3:
    cmd += ["-k", "regex:" + sys.argv[3], "-c", "1"]
txt = subprocess.run(cmd, capture_output=True, text=True).stdout
blocks = txt.split('"Kernel Name",')
blk = blocks[1]
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
# cumulative samples in blocks of 100 SASS instructions (which warp role spends the time)
if len(sys.argv) > 4:
    step = int(sys.argv[4])
    for s in range(0, len(rows), step):
        seg = rows[s:s + step]
        print(f"#{s:<5} samples {sum(float(x[ci['# Samples']] or 0) for x in seg):>8.0f}  exec {sum(float(x[ci['Instructions Executed']] or 0) for x in seg):>12.0f}")
if len(sys.argv) > 6:
    lo, hi = int(sys.argv[5]), int(sys.argv[6])
    seg = rows[lo:hi]
    agg = {k: sum(float(x[ci[k]] or 0) for x in seg) for k in stalls}
    print("range", lo, hi, {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]})
    ops = {}
    for x in seg:
        op = x[ci["Source"]].strip().split()
        op = op[1] if op and op[0].startswith("@") else (op[0] if op else "")
        op = op.split(".")[0]
        o = ops.setdefault(op, [0, 0])
        o[0] += float(x[ci["# Samples"]] or 0); o[1] += float(x[ci["Instructions Executed"]] or 0)
    for k, v in sorted(ops.items(), key=lambda kv: -kv[1][0])[:25]:
        print(f"  {k:<12} samples {v[0]:>8.0f} exec {v[1]:>12.0f}")
if len(sys.argv) > 7 and sys.argv[7] == "list":
    for i in range(lo, hi):
        x = rows[i]
        top = sorted(((float(x[ci[k]] or 0), k) for k in stalls), reverse=True)[:2]
        print(f'#{i:<5}{x[ci["# Samples"]]:>6} ex={x[ci["Instructions Executed"]]:>8} {x[ci["Source"]].strip()[:70]:<70} {[(int(a), b[6:]) for a, b in top if a > 0]}')
