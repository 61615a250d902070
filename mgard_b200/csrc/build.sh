#!/bin/bash
# Builds mgard_b200/libmgard_b200.so for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
OUT=../libmgard_b200.so
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
 -Xcompiler -fPIC,-ffp-contract=off,-O2 --expt-relaxed-constexpr -Xptxas -v"
mkdir -p ../build
pids=()
for f in mgx_hierarchy.cpp mgx_highlevel.cpp mgx_refactor.cu mgx_level.cu mgx_quant.cu mgx_codebook.cu mgx_huffman.cu mgx_hier.cu mgx_hier3.cu mgx_noalloc.cu mgx_api.cu; do
  o=../build/${f%.*}.o
  if [ ! -f $o ] || [ $f -nt $o ] || [ mgx_internal.h -nt $o ] || [ mgx_device.cuh -nt $o ] || [ mgx_hier_common.cuh -nt $o ] || [ ../../include/mgard_b200.h -nt $o ] || [ ../../include/mgard_b200_x.h -nt $o ]; then
    ( $NVCC $FLAGS -x cu -c $f -o $o > ../build/${f%.*}.log 2>&1 || { cat ../build/${f%.*}.log; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o $OUT ../build/*.o
echo "built $OUT"
