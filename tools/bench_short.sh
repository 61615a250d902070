#!/bin/bash
# tests + short bench (stage times) at tol 1e-3 and 1e-5
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3)
for t in 1e-3 1e-5; do
python bench.py --no-cpu --no-e2e --no-extra --tol $t 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['value'],1), round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['roofline']['stages_ms'].items()}, {k:round(v,3) for k,v in b['roofline']['decompress_stages_ms'].items()}, round(b['compression_ratio'],2), b['error_bound']['ok'])
"
done
