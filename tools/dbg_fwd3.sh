#!/bin/bash
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_fwd3.py -x -q 2>&1 | tail -5)
for d in 0 1 2 3; do
echo "MGX_F3_DBG=$d"
MGX_F3_DBG=$d timeout 300 python tools/stage_times.py --shape 1025 1025 1025 --steps 3 --warmup 1 2>&1 | tail -1 | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print(b['stages_ms'])"
done 2>&1 | tee gpurun_out/dbg_fwd3.log
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_fwd3I -c 1 -f -o gpurun_out/full_k_fwd3 python tools/stage_times.py --shape 1025 1025 1025 --steps 1 --warmup 0 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
ls -la gpurun_out/*.ncu-rep
