cd "$(dirname "$0")/.."
for e in -17 -11 -10; do
  PE_BENCH_NO_CPU=1 PE_KNN_EPS_LOG2=$e timeout 300 python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('eps 2^$e bench flagged', d['knn_certificate']['flagged_last_timed_step'], 'scan', round(d['kernels_ms_per_step']['knn_l2_scan'],4), 'select', round(d['kernels_ms_per_step']['knn_select'],4))"
  PE_KNN_EPS_LOG2=$e timeout 300 python tools/run_configs.py --config 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('eps 2^$e cfg5 ms_knn', round(d['ms_knn'],4), 'flagged', d.get('knn_flagged'))"
done
