#!/bin/bash
# GPU box routine for the numbers kept under profiles/: bench line, ncu launch
# list of the same command, one --set full capture per hot kernel.
mkdir -p gpurun_out
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2>> gpurun_out/bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra > gpurun_out/ncu.log 2>&1
for k in k_norm k_hier_forward k_codebook k_encode k_decode_multi k_hier_inverse3d; do
  ncu --set full --import-source on --clock-control none -k regex:$k -c 1 -f -o gpurun_out/full_$k \
      python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-extra >> gpurun_out/ncu.log 2>&1
done
for k in k_level_forward k_lpk_mid k_ipk_contig_ws k_ipk_strided_ws k_level_restore_rows; do
  ncu --set full --import-source on --clock-control none -k regex:$k -c 1 -f -o gpurun_out/full_cfg2_$k \
      python tools/stage_times.py --shape 8193 8193 --dtype f64 --s 0 --tol 1e-4 --steps 1 --warmup 0 >> gpurun_out/ncu.log 2>&1
done
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv \
    --log-file gpurun_out/launches_cfg2.csv python tools/stage_times.py --shape 8193 8193 --dtype f64 --s 0 --tol 1e-4 --steps 1 --warmup 0 >> gpurun_out/ncu.log 2>&1

# reports are too big to travel: summarise here, keep the hot-line pages, drop the rest
python tools/ncu_summary.py gpurun_out/ncu_full_summary.md "ncu --set full --clock-control none (1 x B200, bench.py 1025^3 fp32 rel L-inf 1e-3)" \
    gpurun_out/full_k_*.ncu-rep > gpurun_out/dram_traffic.tsv
python tools/ncu_summary.py gpurun_out/ncu_full_summary_cfg2.md "ncu --set full --clock-control none (1 x B200, 8193^2 fp64 L2 1e-4, finest level)" \
    gpurun_out/full_cfg2_*.ncu-rep >> gpurun_out/dram_traffic.tsv
for k in k_hier_forward k_encode; do
  ncu -i gpurun_out/full_$k.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/src_$k.csv.gz
done
rm -f gpurun_out/*.ncu-rep
cat gpurun_out/bench.json | cut -c1-300
