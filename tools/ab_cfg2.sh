#!/bin/bash
# A/B of library build variants on cfg 2 (8193^2 fp64 L2)
cp mgard_b200/libmgard_b200.so /tmp/lib_base.so
for f in /tmp/lib_base.so mgard_b200/libmgard_b200_*.so; do
  tag=$(basename $f .so)
  [ $f != /tmp/lib_base.so ] && cp $f mgard_b200/libmgard_b200.so
  echo -n "$tag: "
  timeout 300 python tools/stage_times.py --shape 8193 8193 --dtype f64 --s 0 --tol 1e-4 --steps 7 --warmup 2 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['compress_ms'],3), round(b['decompress_ms'],3), {k:round(v,3) for k,v in b['stages_ms'].items()})
"
done
cp /tmp/lib_base.so mgard_b200/libmgard_b200.so
