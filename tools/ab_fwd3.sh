#!/bin/bash
# GPU box routine: parity of the TMA forward pass, then A/B stage times against the
# LDGSTS kernel family (MGX_NO_TMA=1) at tol 1e-3 and 1e-5.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_fwd3.py -x -q 2>&1 | tail -15) > gpurun_out/fwd3_tests.log
tail -6 gpurun_out/fwd3_tests.log
for v in ${AB_VARIANTS:-0 1}; do for t in 1e-3 1e-5; do
echo "MGX_NO_TMA=$v tol=$t"
MGX_NO_TMA=$v timeout 300 python bench.py --no-cpu --no-e2e --no-extra --tol $t 2>gpurun_out/bench_err_$v.log | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['value'],1), round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['roofline']['stages_ms'].items()}, round(b['compression_ratio'],2), b['error_bound']['ok'])
"
done; done 2>&1 | tee gpurun_out/ab_fwd3.log
