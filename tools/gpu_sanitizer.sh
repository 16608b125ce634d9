#!/bin/bash
# compute-sanitizer over a reduced selection of the GPU parity tests: memcheck, racecheck, synccheck,
# initcheck.  usage: tools/gpu_sanitizer.sh [N]  (N = 2: also memcheck of the peer-window all-gather under torchrun)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
N=${1:-1}
SEL='(forward_matches and (rgbdnor_16 or manifold-0 or tensor-0)) or knn_l2_edge or knn_l2_matches_cv2 or hamming_matches or binarize or histogram_confusion or bootstrap or render_bit or near_plane or textured or (miner and batch)'
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ $tool = memcheck ] && extra="--leak-check no --report-api-errors no"
  [ $tool = racecheck ] && extra="--racecheck-report analysis"
  timeout 1200 $CS --tool $tool $extra --error-exitcode 77 --print-limit 20 python -m pytest tests -m gpu -q -x -k "$SEL" > $O/sanitizer_$tool.log 2>&1
  rc=$?
  echo "$tool rc=$rc: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' $O/sanitizer_$tool.log | tr '\n' ' ')"
done
if [ "$N" -ge 2 ]; then
  PE_SANITIZE=1 timeout 900 $CS --tool memcheck --leak-check no --report-api-errors no --target-processes all --error-exitcode 77 --print-limit 20 \
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 tests/multigpu_check.py > $O/sanitizer_p2p_n$N.log 2>&1
  echo "p2p memcheck rc=$?: $(grep -E 'ERROR SUMMARY|p2p_check' $O/sanitizer_p2p_n$N.log | tr '\n' ' ' | cut -c1-600)"
fi
