#!/bin/bash
for m in 2 3; do
  echo "MGX_FW_MINB=$m"
  for c in "8193 8193 --dtype f64 --s 0 --tol 1e-4" "8 16 39 39 1001 --dtype f64 --s 1 --tol 1" "513 513 513 --dtype f32 --s 0 --tol 1e-4"; do
    MGX_FW_MINB=$m python tools/stage_times.py --shape $c 2>&1 | tail -1 | cut -c1-330
  done
done
