#!/bin/bash
# Runs on the GPU box: one plain bench run (must exit 0), then only the `ncu --set full` pass of gpu_profile.sh.
# usage: tools/gpu_ncu_full.sh <tag> <kernel regex> <skip> [count]
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TAG=$1; KREGEX=$2; SKIP=$3; COUNT=${4:-4}
BENCH="python bench.py --steps 5 --warmup 3"
if timeout 600 $BENCH > $O/bench_$TAG.log 2> $O/bench_$TAG.err; then
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s $SKIP -c $COUNT -f -o $O/prof_$TAG $BENCH > $O/ncu_full_$TAG.log 2>&1
  echo "ncu full exit $?" >> $O/ncu_full_$TAG.log
  tail -n 2 $O/ncu_full_$TAG.log
else
  echo "bench failed"; tail -n 5 $O/bench_$TAG.err
fi
