#!/bin/bash
# Developer probe: builds librc variants with -D overrides into raytracing_test_b200/variants/ (git-ignored *.so).
# usage: tools/build_variants.sh tag1 "-DA=1 -DB=2" tag2 "..." ...
set -e
cd "$(dirname "$0")/../raytracing_test_b200/csrc"
mkdir -p ../variants
BUILD_ID=$(cat kernels.cu kernels.h free_flight.h | sha256sum | cut -d' ' -f1)
while [ $# -ge 2 ]; do
  tag=$1; defs=$2; shift 2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17 -ccbin /usr/bin/g++ \
    -Xcompiler -fPIC,-ffp-contract=off -DRC_BUILD_ID=\"$BUILD_ID\" $defs -Xptxas -v -shared -o ../variants/librc_$tag.so rc_api.cu kernels.cu worldgen.cu flatten.cpp world_blob.cpp 2> ../variants/build_$tag.log &
done
wait
grep -A2 "primary_kernelE" ../variants/build_*.log | grep -E "Used" || true
