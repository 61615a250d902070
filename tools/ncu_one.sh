#!/bin/bash
# usage: ncu_one.sh <kernel regex> <out name> -- one --set full capture of the bench workload
# (NCU_ARGS: extra arguments of tools/stage_times.py, default the bench's NYX-shaped field)
mkdir -p gpurun_out
python tools/stage_times.py --shape 1025 1025 1025 --steps 1 --warmup 0 ${NCU_ARGS:---field nyx} > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --import-source on --clock-control none -k regex:$1 -c ${NCU_COUNT:-1} -f -o gpurun_out/$2 python tools/stage_times.py --shape 1025 1025 1025 --steps 1 --warmup 0 ${NCU_ARGS:---field nyx} > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
ncu -i gpurun_out/$2.ncu-rep --page raw --csv > gpurun_out/$2_raw.csv 2>/dev/null
ncu -i gpurun_out/$2.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/$2_src.csv.gz
ls -la gpurun_out/
