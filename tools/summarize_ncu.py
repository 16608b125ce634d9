#!/usr/bin/env python
"""Turns the two ncu passes of tools/gpu_profile.sh into the small text summaries committed under profiles/.

  python tools/summarize_ncu.py <tag> [--steps 5 --warmup 3 --launches-per-step 6]

  gpurun_out/launches_<tag>.csv  -> profiles/<tag>_launches.md   (per-kernel share of the timed steps)
  gpurun_out/prof_<tag>.ncu-rep  -> profiles/<tag>_ncu_full.md   (dram bytes, throughput %, registers ... per capture)
                                    profiles/ncu_traffic.json     (dram bytes per launch of the dominant kernels)
"""
import argparse
import csv
import io
import json
import os
import re
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.avg",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_bytes.sum", "smsp__cycles_active.avg",
    "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__shared_mem_per_block_dynamic",
]


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.replace("pe::<unnamed>::", "").replace("unnamed>::", "").replace("void ", "").strip()


def launches(tag, steps, warmup, per_step):
    path = os.path.join(ROOT, "gpurun_out", f"launches_{tag}.csv")
    rows = []
    import gzip
    opener = (lambda: gzip.open(path + ".gz", "rt")) if os.path.exists(path + ".gz") else (lambda: open(path))
    with opener() as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(io.StringIO("".join(lines))):
        if r["Metric Name"] == "gpu__time_duration.sum":
            rows.append((int(r["ID"]), short(r["Kernel Name"]), float(r["Metric Value"]) / 1e3, r["Grid Size"], r["Block Size"]))
    # the timed region of bench.py: `steps` steps that follow the set-up and the warm-up.  Locate it as the
    # last steps*per_step launches before the first launch of the host-API (e2e) phase, which is the first
    # conv1 launch with a grid smaller than a full batch (1024-patch H2D stages) -- or fall back to the tail.
    names = [r[1] for r in rows]
    first_scan = next(i for i, n in enumerate(names) if "scan" in n)
    start = first_scan - 4 + warmup * per_step          # first launch of timed step 0 (conv1, conv2, fc, fc precede the scan)
    timed = rows[start:start + steps * per_step]
    agg = OrderedDict()
    for _, n, us, g, b in timed:
        a = agg.setdefault(n, [0, 0.0, g, b])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    out = [f"# ncu launch list, bench.py --steps {steps} --warmup {warmup} ({tag})", "",
           "`ncu --metrics gpu__time_duration.sum --clock-control none` over the whole bench command; the table is the",
           f"timed region only (launches {timed[0][0]}..{timed[-1][0]} of {len(rows)}).  Times are cold-cache and serialised:",
           "compare the SHARES with bench.py's CUDA-event numbers, not the absolutes.", "",
           "| kernel | launches | total us | us / launch | share | grid | block |", "|---|---|---|---|---|---|---|"]
    for n, (c, us, g, b) in agg.items():
        out.append(f"| `{n}` | {c} | {us:.1f} | {us / c:.1f} | {100 * us / total:.1f} % | {g} | {b} |")
    out.append(f"| total | {len(timed)} | {total:.1f} | | | | |")
    allagg = OrderedDict()
    for _, n, us, g, b in rows:
        a = allagg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += us
    out += ["", "Whole process (set-up of the 1,000,200-row DB, warm-up, timed steps, e4m3 / serialised / sustained replays, host-API e2e steps):", "",
            "| kernel | launches | total us |", "|---|---|---|"]
    for n, (c, us) in allagg.items():
        out.append(f"| `{n}` | {c} | {us:.1f} |")
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", f"{tag}_launches.md"), "w") as f:
        f.write("\n".join(out) + "\n")
    return {n: us / c for n, (c, us, _, _) in agg.items()}


def full(tag):
    rep = os.path.join(ROOT, "gpurun_out", f"prof_{tag}.ncu-rep")
    if not os.path.exists(rep):
        return {}
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rd[0], rd[1], rd[2:]
    col = {h: i for i, h in enumerate(hdr)}
    out = [f"# ncu --set full captures ({tag})", "",
           "`ncu --set full --clock-control none --import-source on` on `python bench.py --steps 5 --warmup 3`; raw page, selected metrics.", ""]
    traffic = {}
    for r in data:
        name = short(r[col["Kernel Name"]])
        out.append(f"## `{name}`  (launch id {r[col['ID']]}, grid {r[col['Grid Size']]}, block {r[col['Block Size']]})")
        out.append("")
        out.append("| metric | value | unit |")
        out.append("|---|---|---|")
        for k in KEYS:
            if k in col:
                out.append(f"| {k} | {r[col[k]]} | {units[col[k]]} |")
        out.append("")

        def num(k):
            try:
                v = float(r[col[k]].replace(",", ""))
            except Exception:
                return None
            u = units[col[k]].lower()
            mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            return v * mult
        rd_b, wr_b = num("dram__bytes_read.sum"), num("dram__bytes_write.sum")
        if rd_b is not None and wr_b is not None:
            traffic.setdefault(name, []).append(rd_b + wr_b)
    with open(os.path.join(ROOT, "profiles", f"{tag}_ncu_full.md"), "w") as f:
        f.write("\n".join(out) + "\n")
    return {k: sum(v) / len(v) for k, v in traffic.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("tag")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    # round 2: conv1, conv2, fc x2, scan, select, exact re-scan of unproven queries (knn_exact_fallback: exits at once when none)
    ap.add_argument("--launches-per-step", type=int, default=7)
    a = ap.parse_args()
    per = launches(a.tag, a.steps, a.warmup, a.launches_per_step)
    tr = full(a.tag)
    tj = {"tag": a.tag, "dram_bytes_per_launch": tr, "ncu_us_per_launch_timed_region": per}
    for k, v in tr.items():
        if "conv1_tc" in k:
            tj["conv1_tc_bytes_per_launch"] = v
        elif "conv1_pool_act" in k:
            tj["conv1_pool_act_bytes_per_launch"] = v
        if "scan" in k:
            tj["knn_scan_bytes_per_launch"] = v
    with open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w") as f:
        json.dump(tj, f, indent=1)
    print(json.dumps(tj, indent=1))


if __name__ == "__main__":
    sys.exit(main())
