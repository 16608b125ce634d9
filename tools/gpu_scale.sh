#!/bin/bash
# Runs on an N-GPU box: bench.py at N ranks, then BASELINE configs 3 and 5 sharded over N ranks.  usage: gpu_scale.sh N
cd "$(dirname "$0")/.."
N=$1; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.log 2> $O/bench_n$N.err; echo "bench rc=$?"
for c in 3 5; do
  timeout 600 $TR --master-port 2953$c tools/run_configs.py --config $c > $O/cfg${c}_n$N.log 2> $O/cfg${c}_n$N.err; echo "config $c rc=$?"
done
python - <<PY
import json
for f in ("bench_n$N","cfg3_n$N","cfg5_n$N"):
    try:
        d=json.loads(open("$O/"+f+".log").read().strip().splitlines()[-1])
        keep={k:d[k] for k in d if k in ("value","ms_per_step","n_gpus","descriptors_per_s","knn_queries_per_s","patches_per_s_forward_plus_knn","ms_forward","ms_knn","ms_step","clocks","kernels_ms_per_step")}
        if "e2e" in d: keep["e2e"]=d["e2e"]["value"]
        print(f, json.dumps(keep))
    except Exception as e:
        print(f, "failed", e)
PY
