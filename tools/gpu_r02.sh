#!/bin/bash
# Round-2 GPU run: parity tests, smoke, bench; optionally the two ncu passes.  usage: tools/gpu_r02.sh <tag> [ncu] [pytest -k expr]
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TAG=${1:-r02a}; NCU=${2:-no}; KEXPR=${3:-}
nvidia-smi -L > $O/smi_$TAG.txt
if [ -n "$KEXPR" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -x -k "$KEXPR" > $O/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_$TAG.log
else
  timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_$TAG.log
fi
tail -n 15 $O/pytest_$TAG.log
timeout 300 python __graft_entry__.py smoke > $O/smoke_$TAG.log 2>&1; echo "smoke exit $?" >> $O/smoke_$TAG.log
tail -n 2 $O/smoke_$TAG.log
BENCH="python bench.py --steps 10 --warmup 3"
if timeout 900 $BENCH > $O/bench_$TAG.log 2> $O/bench_$TAG.err; then
  python - <<PY
import json
d=json.loads(open("$O/bench_$TAG.log").read().strip().splitlines()[-1])
print("value",round(d["value"]),"ms/step",round(d["ms_per_step"],4),"serial",round(d["ms_per_step_serialised"],4),"e4m3",round(d["value_e4m3_optin"]),"e2e",round(d["e2e"]["value"]),"sus",round(d["sustained"]["ms_per_step"],4))
print({k:round(v,4) for k,v in d["kernels_ms_per_step"].items()})
print(d["parity_in_run"])
print("roofline frac",round(d["roofline"]["frac"],4),"cpu",d.get("cpu_baseline",{}).get("value"))
PY
  if [ "$NCU" = "ncu" ]; then
    BQ="python bench.py --steps 5 --warmup 3"
    timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_$TAG.csv $BQ > $O/ncu_launches_$TAG.log 2>&1
    echo "ncu launches exit $?" >> $O/ncu_launches_$TAG.log
    timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:conv1_tc|knn_tc_scan" -s 260 -c 4 -f -o $O/prof_$TAG $BQ > $O/ncu_full_$TAG.log 2>&1
    echo "ncu full exit $?" >> $O/ncu_full_$TAG.log
    tail -n 2 $O/ncu_launches_$TAG.log $O/ncu_full_$TAG.log
  fi
else
  echo "bench failed"; tail -n 30 $O/bench_$TAG.err
fi
