#!/bin/bash
# usage: ncu_cfg.sh <kernel regex> <out name> <count> <stage_times args...>: --set full capture of the first <count> matches
re=$1; name=$2; cnt=$3; shift 3
mkdir -p gpurun_out
python tools/stage_times.py --steps 1 --warmup 0 "$@" > gpurun_out/plain_$name.log 2>&1 &&
timeout 900 ncu --set full --import-source on --clock-control none -k regex:$re -c $cnt -f -o gpurun_out/$name python tools/stage_times.py --steps 1 --warmup 0 "$@" > gpurun_out/ncu_$name.log 2>&1
tail -2 gpurun_out/ncu_$name.log
ncu -i gpurun_out/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
ncu -i gpurun_out/$name.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/${name}_src.csv.gz
