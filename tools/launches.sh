#!/bin/bash
# launch list (cold-cache, serialised: compare shares) of one compress + decompress of the bench workload
mkdir -p gpurun_out
python tools/stage_times.py --shape 1025 1025 1025 --steps 1 --warmup 1 "$@" > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 200 --csv --log-file gpurun_out/launches.csv python tools/stage_times.py --shape 1025 1025 1025 --steps 1 --warmup 1 "$@" > gpurun_out/ncu.log 2>&1
python tools/launch_summary.py gpurun_out/launches.csv | tail -30
