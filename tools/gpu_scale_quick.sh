#!/bin/bash
# bench.py + tests/multigpu_check.py at N ranks only (no configs, no reference arm).  usage: gpu_scale_quick.sh N
cd "$(dirname "$0")/.."
N=$1; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tests/multigpu_check.py > $O/p2p_n$N.log 2> $O/p2p_n$N.err; echo "multigpu_check rc=$?"; tail -n 1 $O/p2p_n$N.log | cut -c1-400
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.log 2> $O/bench_n$N.err; echo "bench rc=$?"
tail -n 1 $O/bench_n$N.log | python -c "
import json,sys; d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], 'value', round(d['value']), 'ms', round(d['ms_per_step'],4), 'serial', round(d['ms_per_step_serialised'],4), 'coll', round(d['collective_ms_per_step'],4), 'e2e', round(d['e2e']['value']), {k:round(v,4) for k,v in d['kernels_ms_per_step'].items()}, d['parity_in_run']['sharded_equals_unsharded'], d['parity_in_run']['descriptor_rowscale_err'])"
