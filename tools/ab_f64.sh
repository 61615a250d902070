#!/bin/bash
# A/B of library build variants on fp64 3-D fields (k_fwd3<double>)
cp mgard_b200/libmgard_b200.so /tmp/lib_base.so
for f in /tmp/lib_base.so mgard_b200/libmgard_b200_*.so; do
  tag=$(basename $f .so)
  [ $f != /tmp/lib_base.so ] && cp $f mgard_b200/libmgard_b200.so
  for fld in nyx smooth; do
    echo -n "$tag f64 641^3 $fld: "
    timeout 300 python tools/stage_times.py --shape 641 641 641 --dtype f64 --steps 7 --warmup 2 --tol 1e-3 --field $fld 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['compress_ms'],3), round(b['decompress_ms'],3), {k:round(v,3) for k,v in b['stages_ms'].items()})
"
  done
done
cp /tmp/lib_base.so mgard_b200/libmgard_b200.so
