#!/bin/bash
# A/B of conv1's issue order (see conv_tc.cu): default = cross terms first, PE_C1_DBG=512 = the three products of a K step
# back to back (round 1).  Prints conv1 ms and the descriptor error of the timed region for both.
cd "$(dirname "$0")/.."
for o in 0 512; do
  PE_BENCH_NO_CPU=1 PE_C1_DBG=$o timeout 300 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('PE_C1_DBG=$o conv1', round(d['kernels_ms_per_step']['conv1_pool_act'],4), 'step', round(d['ms_per_step'],4), 'rowscale', '%.2e'%d['parity_in_run']['descriptor_rowscale_err'], 'elemfrac', round(d['parity_in_run']['descriptor_elementrel_frac_within_1e-4'],4))"
done
