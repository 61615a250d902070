#!/bin/bash
# GPU box routine: parity tests, then stage timings of the correction-path configs.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/gpu_tests.log
tail -4 gpurun_out/gpu_tests.log
python tools/stage_times.py --shape 8193 8193 --dtype f64 --s 0 --tol 1e-4 2>&1 | tail -1 | tee gpurun_out/cfg2.json
python tools/stage_times.py --shape 8 16 39 39 1001 --dtype f64 --s 1 --tol 1 2>&1 | tail -1 | tee gpurun_out/cfg4s.json
python tools/stage_times.py --shape 513 513 513 --dtype f32 --s 0 --tol 1e-4 2>&1 | tail -1 | tee gpurun_out/cfg_3d_l2.json
