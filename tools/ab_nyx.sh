#!/bin/bash
# A/B of library build variants on the NYX-shaped field at several tolerances
# (tools/stage_times.py --field nyx); AB_ENV="VAR=1" adds a run of the base build with that env
cp mgard_b200/libmgard_b200.so /tmp/lib_base.so
run() { # tag tol field [env]
  echo -n "$1 $3 tol=$2 $4: "
  env $4 timeout 300 python tools/stage_times.py --shape 1025 1025 1025 --steps 5 --warmup 2 --tol $2 --field $3 $AB_ARGS 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(b['compress_ms'],3), round(b['decompress_ms'],3), {k:round(v,3) for k,v in b['stages_ms'].items()}, round(b['CR'],2), round(b['err_over_tol'],3))
"
}
for f in /tmp/lib_base.so mgard_b200/libmgard_b200_*.so; do
  tag=$(basename $f .so)
  [ $f != /tmp/lib_base.so ] && cp $f mgard_b200/libmgard_b200.so
  for t in ${AB_TOLS:-1e-3 1e-5}; do
    for fld in ${AB_FIELDS:-nyx}; do run $tag $t $fld; done
  done
done
cp /tmp/lib_base.so mgard_b200/libmgard_b200.so
if [ -n "$AB_ENV" ]; then for t in ${AB_TOLS:-1e-3 1e-5}; do run base $t nyx "$AB_ENV"; done; fi
