#!/bin/bash
# BASELINE configs 1, 3, 4, 5, 6 on ONE GPU (tools/run_configs.py) + the 8f widening benches; raw JSON lines into gpurun_out/
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for c in 1 3 4 5 6; do
  timeout 600 python tools/run_configs.py --config $c > $O/cfg${c}_n1.log 2> $O/cfg${c}_n1.err; echo "config $c rc=$?"; tail -n 1 $O/cfg${c}_n1.log | cut -c1-600
done
for b in patch render miner; do
  timeout 600 python tools/${b}_bench.py > $O/${b}_bench.log 2> $O/${b}_bench.err; echo "$b rc=$?"; tail -n 1 $O/${b}_bench.log | cut -c1-700
done
