#!/bin/bash
# A/B of library build variants over several shapes (smooth field): "nz ny nx tol" per entry of AB_CASES
cp mgard_b200/libmgard_b200.so /tmp/lib_base.so
IFS=';' read -ra CASES <<< "${AB_CASES:-513 513 513 1e-3;257 2049 2049 1e-4;1025 1025 1025 1e-3}"
for f in /tmp/lib_base.so mgard_b200/libmgard_b200_*.so; do
  tag=$(basename $f .so)
  [ $f != /tmp/lib_base.so ] && cp $f mgard_b200/libmgard_b200.so
  for c in "${CASES[@]}"; do
    set -- $c
    echo -n "$tag $1x$2x$3 tol=$4: "
    timeout 300 python tools/stage_times.py --shape $1 $2 $3 --steps 7 --warmup 2 --tol $4 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['compress_ms'],3), round(b['decompress_ms'],3), {k:round(v,3) for k,v in b['stages_ms'].items()})
"
  done
done
cp /tmp/lib_base.so mgard_b200/libmgard_b200.so
