#!/bin/bash
# Builds library variants (different -D flags for rtb_trace.cu / rtb_api.cu) as variants/librtb200_<name>.so for A/B timing.
# usage: scripts/build_variants.sh name1:"-DFOO=1" name2:"-DFOO=2" ...
set -e
cd "$(dirname "$0")/../ray-tracer-rust_b200"
mkdir -p ../variants build
NV="nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-ffp-contract=off"
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  $NV $flags -Xptxas -v -c csrc/rtb_trace.cu -o build/rtb_trace_$name.o 2> build/rtb_trace_$name.log &
  $NV $flags -c csrc/rtb_api.cu -o build/rtb_api_$name.o &
  $NV $flags -c csrc/rtb_grid.cu -o build/rtb_grid_$name.o &
  wait
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/librtb200_$name.so build/rtb_api_$name.o build/rtb_trace_$name.o build/rtb_grid_$name.o build/rtb_bvh.o build/rtb_refpair.o build/rtb_peaks.o build/rtb_host.o build/rtb_obj_parallel.o -cudart shared
  $NV $flags -Xptxas -v -c csrc/rtb_grid.cu -o /dev/null 2> build/rtb_grid_$name.log || true
  echo "$name: $(grep -A2 'k_shadow_gridILb0\|k_primaryILb0\|k_shadeILb0ELb0' build/rtb_trace_$name.log build/rtb_grid_$name.log | grep -o 'Used [0-9]* registers[^,]*, [0-9]* bytes smem\|Used [0-9]* registers' | tr '\n' ' ')"
done
