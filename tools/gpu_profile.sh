#!/bin/bash
# Runs on the GPU box (via gpurun): GPU parity tests, smoke, bench, then the two ncu passes of
# /opt/skills/guides/B200_PROFILING.md on the same bench command.  Outputs land in gpurun_out/.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TAG=${1:-r01}
KREGEX=${2:-'conv1_pool_act|knn_l2_scan'}
SKIP=${3:-248}
BENCH="python bench.py --steps 5 --warmup 3"
nvidia-smi -L > $O/smi.txt
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu_$TAG.log
timeout 300 python __graft_entry__.py smoke > $O/smoke_$TAG.log 2>&1; echo "smoke exit $?" >> $O/smoke_$TAG.log
if timeout 600 $BENCH > $O/bench_$TAG.log 2> $O/bench_$TAG.err; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_$TAG.csv $BENCH > $O/ncu_launches_$TAG.log 2>&1
  echo "ncu launches exit $?" >> $O/ncu_launches_$TAG.log
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s $SKIP -c 4 -f -o $O/prof_$TAG $BENCH > $O/ncu_full_$TAG.log 2>&1
  echo "ncu full exit $?" >> $O/ncu_full_$TAG.log
else
  echo "bench failed" >> $O/bench_$TAG.err
fi
tail -n 3 $O/pytest_gpu_$TAG.log $O/smoke_$TAG.log $O/bench_$TAG.err
