#!/bin/bash
# A/B of library build variants with tools/stage_times.py (all stage times of compress and decompress)
cp mgard_b200/libmgard_b200.so /tmp/lib_base.so
for f in /tmp/lib_base.so mgard_b200/libmgard_b200_*.so; do
  tag=$(basename $f .so)
  [ $f != /tmp/lib_base.so ] && cp $f mgard_b200/libmgard_b200.so
  for t in ${AB_TOLS:-1e-3}; do
  echo -n "$tag tol=$t: "
  timeout 300 python tools/stage_times.py --shape 1025 1025 1025 --steps 5 --warmup 2 --tol $t 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['compress_ms'],3), round(b['decompress_ms'],3), {k:round(v,3) for k,v in b['stages_ms'].items()}, round(b['CR'],2), round(b['err_over_tol'],3))
"
  done
done
cp /tmp/lib_base.so mgard_b200/libmgard_b200.so
