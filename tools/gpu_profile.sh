#!/bin/bash
# Runs on the GPU box (via gpurun): one plain bench run (must exit 0), then the two ncu passes of
# /opt/skills/guides/B200_PROFILING.md on the same bench command.  Outputs land in gpurun_out/.
# usage: tools/gpu_profile.sh <tag> [kernel regex] [skip]
#   skip: matching launches before the timed region = 245 set-up conv1 launches + 3 warm-up steps x (conv1 + scan) = 251
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TAG=${1:-r02a}
KREGEX=${2:-'conv1_tc|knn_tc_scan'}
SKIP=${3:-252}
BENCH="python bench.py --steps 5 --warmup 3"
nvidia-smi -L > $O/smi_$TAG.txt
if timeout 900 $BENCH > $O/bench_$TAG.log 2> $O/bench_$TAG.err; then
  export PE_BENCH_NO_CPU=1      # the two profiler passes skip the CPU-baseline leg (host-only; nothing to profile there)
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_$TAG.csv $BENCH > $O/ncu_launches_$TAG.log 2>&1
  echo "ncu launches exit $?" >> $O/ncu_launches_$TAG.log
  timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s $SKIP -c 4 -f -o $O/prof_$TAG $BENCH > $O/ncu_full_$TAG.log 2>&1
  echo "ncu full exit $?" >> $O/ncu_full_$TAG.log
  gzip -f $O/launches_$TAG.csv
  tail -n 1 $O/bench_$TAG.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms/step',round(d['ms_per_step'],4),'launches',d['gpu_launches'],'kernels',{k:round(v,4) for k,v in d['kernels_ms_per_step'].items()})"
  tail -n 2 $O/ncu_launches_$TAG.log $O/ncu_full_$TAG.log
else
  echo "bench failed"; tail -n 20 $O/bench_$TAG.err
fi
