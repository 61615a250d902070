#!/bin/bash
# GPU box routine for the numbers kept under profiles/ (one ncu invocation per gpurun
# call on this pool -- run the pieces separately, see the comments):
#   bash tools/gpu_profile.sh bench     bench lines (own arm + reference arm) + launch list
#   bash tools/gpu_profile.sh cfg3      --set full of the hot kernels, 1025^3 fp32 1e-3
#   bash tools/gpu_profile.sh cfg2      --set full of the correction-path kernels, 8193^2 fp64
# then, back in the container:
#   python tools/ncu_summary.py profiles/r02_ncu_full_summary.md "<title>" gpurun_out/r02_full.ncu-rep
mkdir -p gpurun_out
case "$1" in
bench)
  python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || exit 1
  python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
  python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra > gpurun_out/bench_short.json 2>/dev/null &&
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv \
      --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra > gpurun_out/ncu_bench.log 2>&1
  ;;
cfg3)
  NCU_COUNT=24 bash tools/ncu_one.sh "k_fwd3|k_encode|k_inv_level3|k_decode_multi|k_norm|k_codebook|k_hist_fix" r02_full
  ;;
cfg2)
  bash tools/ncu_cfg.sh "k_level_forward|k_lpk_mid|k_ipk_part|k_level_restore_rows|k_lpk_first" r02_cfg2 8 \
      --shape 8193 8193 --dtype f64 --s 0 --tol 1e-4
  ;;
*)
  echo "usage: gpu_profile.sh bench|cfg3|cfg2"; exit 2 ;;
esac
