#!/bin/bash
# ncu --set full of the step's smaller kernels (conv2, the two InnerProduct launches, the select) on the bench command,
# after a plain run that must exit 0.  usage: tools/gpu_profile_small.sh <tag>
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TAG=${1:-r02c}
BENCH="python bench.py --steps 5 --warmup 3"
export PE_BENCH_NO_CPU=1
if timeout 600 $BENCH > $O/bench_small_$TAG.log 2> $O/bench_small_$TAG.err; then
  # matching launches before the timed region: 245 set-up batches x (conv2 + 2 fc) + 3 warm-up steps x 4
  timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:conv2_tc|fc_tc|knn_tc_select" -s 752 -c 8 -f -o $O/prof_small_$TAG $BENCH > $O/ncu_small_$TAG.log 2>&1
  echo "ncu exit $?" >> $O/ncu_small_$TAG.log
  tail -n 3 $O/ncu_small_$TAG.log
  ncu -i $O/prof_small_$TAG.ncu-rep --page raw --csv > $O/prof_small_$TAG.csv 2>/dev/null; wc -l $O/prof_small_$TAG.csv
else
  echo "bench failed"; tail -n 20 $O/bench_small_$TAG.err
fi
