#!/bin/bash
# usage: launches_cfg.sh <name> <stage_times args...>: launch list (cold-cache, serialised) of ONE compress + decompress
name=$1; shift
mkdir -p gpurun_out
python tools/stage_times.py --steps 1 --warmup 0 "$@" > gpurun_out/plain_$name.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$name.csv python tools/stage_times.py --steps 1 --warmup 0 "$@" > gpurun_out/ncu_$name.log 2>&1
tail -2 gpurun_out/ncu_$name.log
