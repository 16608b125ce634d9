#!/bin/bash
# quick GPU check: selected parity tests + a short bench; prints a one-line summary
cd "$(dirname "$0")/.."
TAG=${1:-q}; KEXPR=${2:-"knn or full_size_db or sharded"}
timeout 900 python -m pytest tests -m gpu -q -k "$KEXPR" > gpurun_out/pytest_$TAG.log 2>&1; echo "exit $?" >> gpurun_out/pytest_$TAG.log
tail -4 gpurun_out/pytest_$TAG.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_$TAG.log 2> gpurun_out/bench_$TAG.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_$TAG.log").read().strip().splitlines()[-1])
    print("value",round(d["value"]),"ms/step",round(d["ms_per_step"],3),"e2e",round(d["e2e"]["value"]),"e2e ms",round(d["e2e"]["ms_per_step"],2))
    print({k:round(v,4) for k,v in d["kernels_ms_per_step"].items()})
except Exception as e:
    print("bench parse failed",e); print(open("gpurun_out/bench_$TAG.err").read()[-2000:])
PY
