#!/usr/bin/env python
"""SASS opcode histogram of the tensor kernels in the in-tree libpose_b200.so (evidence that the tcgen05 / TMEM / bulk-copy
path is what runs): counts of UTCHMMA / UTCQMMA (tcgen05.mma kind::f16 / kind::f8f6f4), LDTM (tcgen05.ld), UTCBAR
(tcgen05.commit), UBLKCP (cp.async.bulk), SYNCS (mbarrier), UTCATOMSWS / UTCCP (TMEM allocation) per kernel.
  python tools/sass_histogram.py > profiles/r02_sass.md"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "3dposeestimation_b200", "libpose_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "UTCBAR", "UBLKCP", "SYNCS", "UTCATOMSWS", "ELECT", "HMMA", "FFMA", "POPC"]
kern, rows = None, collections.OrderedDict()
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = kern.replace("(anonymous namespace)::", "").replace("void ", "")
        kern = re.sub(r"\(.*", "", kern).replace("pe::", "")
        rows.setdefault(kern, collections.Counter())
        continue
    if kern:
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            rows[kern]["total"] += 1
            for o in OPS:
                if op.startswith(o):
                    rows[kern][o] += 1
print("# SASS opcode histogram of `3dposeestimation_b200/libpose_b200.so` (sm_100a), `tools/sass_histogram.py`\n")
print("`UTCHMMA` = tcgen05.mma kind::f16, `UTCQMMA` = kind::f8f6f4, `LDTM` = tcgen05.ld, `UTCBAR` = tcgen05.commit, `UBLKCP` = cp.async.bulk,")
print("`SYNCS` = mbarrier ops, `UTCATOMSWS` = TMEM alloc / dealloc.  Kernels with any tensor / bulk-copy opcode first, then the rest.\n")
print("| kernel | instr | " + " | ".join(OPS) + " |")
print("|---|---|" + "---|" * len(OPS))
def line(k, c):
    return f"| `{k}` | {c['total']} | " + " | ".join(str(c[o]) if c[o] else "" for o in OPS) + " |"
tens = [(k, c) for k, c in rows.items() if c["UTCHMMA"] or c["UTCQMMA"] or c["UBLKCP"]]
rest = [(k, c) for k, c in rows.items() if not (c["UTCHMMA"] or c["UTCQMMA"] or c["UBLKCP"])]
for k, c in tens + rest:
    print(line(k, c))
