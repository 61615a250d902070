#!/bin/bash
# usage: build_variant.sh <tag> <source file in csrc> <-D...>  -> mgard_b200/libmgard_b200_<tag>.so
# (one translation unit rebuilt with extra defines, linked with the product's other objects)
set -e
tag=$1; src=$2; shift 2
cd "$(dirname "$0")/../mgard_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-ffp-contract=off,-O2 --expt-relaxed-constexpr"
mkdir -p ../build/var
$NVCC $FLAGS "$@" -x cu -c $src -o ../build/var/${tag}.o
objs=$(ls ../build/*.o | grep -v "/${src%.*}.o")
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libmgard_b200_${tag}.so $objs ../build/var/${tag}.o
echo built libmgard_b200_${tag}.so
