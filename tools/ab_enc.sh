#!/bin/bash
# parity, then A/B of the chain-free and look-back encoders at tol 1e-3 / 1e-5
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fwd3.py tests/test_golden.py -x -q 2>&1 | tail -3)
for v in 0 1; do for t in 1e-3 1e-5; do
echo -n "MGX_ENC_LOOKBACK=$v tol=$t: "
MGX_ENC_LOOKBACK=$v timeout 300 python bench.py --no-cpu --no-e2e --no-extra --tol $t 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['value'],1), round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['roofline']['stages_ms'].items()}, round(b['compression_ratio'],2), b['error_bound']['ok'])
"
done; done
