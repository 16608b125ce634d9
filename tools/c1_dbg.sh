#!/bin/bash
# development only: conv1 ablations through PE_C1_DBG (see C1Args::dbg); results of flagged runs are wrong by design
for d in ${@:-0 1 2 3 4 16 20 8 23}; do
  PE_C1_DBG=$d timeout 200 python bench.py --steps 5 --warmup 3 > /tmp/dbg_$d.log 2> /tmp/dbg_$d.err
  python - <<PY
import json
try:
    d=json.loads(open("/tmp/dbg_$d.log").read().strip().splitlines()[-1]); print("dbg=$d conv1 ms", round(d["kernels_ms_per_step"]["conv1_pool_act"],4))
except Exception as e:
    print("dbg=$d failed:", open("/tmp/dbg_$d.err").read()[-300:].replace("\n"," | "))
PY
done
