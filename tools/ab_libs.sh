#!/bin/bash
# A/B of library build variants: mgard_b200/libmgard_b200_<tag>.so copied over the product .so in turn
cp mgard_b200/libmgard_b200.so /tmp/lib_base.so
for f in mgard_b200/libmgard_b200_*.so; do
  tag=$(basename $f .so)
  cp $f mgard_b200/libmgard_b200.so
  for t in 1e-3 ; do
  echo -n "$tag tol=$t: "
  timeout 300 python bench.py --no-cpu --no-e2e --no-extra --tol $t 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['value'],1), {k:round(v,3) for k,v in b['roofline']['stages_ms'].items()}, round(b['compression_ratio'],2), b['error_bound']['ok'])
"
  done
done
cp /tmp/lib_base.so mgard_b200/libmgard_b200.so
