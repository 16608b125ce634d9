#!/bin/bash
# Runs on an N-GPU box: multi-GPU parity (tests/multigpu_check.py), bench.py at N ranks (both arms), BASELINE configs 3, 4, 5, 6
# sharded over N ranks.  Raw JSON lines land in gpurun_out/ (copied to profiles/ for the record).  usage: gpu_scale.sh N [quick]
cd "$(dirname "$0")/.."
N=$1; QUICK=${2:-}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/topo_n$N.txt 2>&1
timeout 600 $TR --master-port 29511 tests/multigpu_check.py > $O/p2p_n$N.log 2> $O/p2p_n$N.err; echo "p2p_check rc=$?"; tail -n 1 $O/p2p_n$N.log
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.log 2> $O/bench_n$N.err; echo "bench rc=$?"
for c in 3 4 5 6; do
  timeout 600 $TR --master-port 2953$c tools/run_configs.py --config $c > $O/cfg${c}_n$N.log 2> $O/cfg${c}_n$N.err; echo "config $c rc=$?"
done
if [ -z "$QUICK" ]; then
  timeout 600 $TR --master-port 29541 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/benchref_n$N.log 2> $O/benchref_n$N.err; echo "reference arm rc=$?"
fi
python - <<PY
import json
for f in ("bench_n$N","cfg3_n$N","cfg4_n$N","cfg5_n$N","cfg6_n$N"):
    try:
        d=json.loads(open("$O/"+f+".log").read().strip().splitlines()[-1])
        keep={k:d[k] for k in d if k in ("value","ms_per_step","ms_per_step_serialised","collective_ms_per_step","n_gpus","descriptors_per_s","knn_queries_per_s","patches_per_s_forward_plus_knn","ms_forward","ms_knn","ms_step","ms_histogram","histogram","kernels_ms_per_step")}
        if "e2e" in d: keep["e2e"]=d["e2e"]["value"]
        if "parity_in_run" in d: keep["sharded_equals_unsharded"]=d["parity_in_run"]["sharded_equals_unsharded"]
        print(f, json.dumps(keep))
    except Exception as e:
        print(f, "failed", e); print(open("$O/"+f+".err").read()[-1500:])
PY
