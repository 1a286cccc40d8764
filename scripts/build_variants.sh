#!/bin/bash
# Builds kernel variants of rtb_trace.cu (different -D flags) as gpurun_variants/librtb200_<name>.so for A/B timing.
# usage: scripts/build_variants.sh name1:"-DFOO=1" name2:"-DFOO=2" ...
set -e
cd "$(dirname "$0")/../ray-tracer-rust_b200"
mkdir -p ../variants build
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-ffp-contract=off $flags -Xptxas -v \
       -c csrc/rtb_trace.cu -o build/rtb_trace_$name.o 2> build/rtb_trace_$name.log
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/librtb200_$name.so build/rtb_api.o build/rtb_trace_$name.o build/rtb_bvh.o build/rtb_refpair.o build/rtb_host.o -cudart shared
  echo "$name: $(grep -A2 'k_trace_frameILb0' build/rtb_trace_$name.log | grep -o 'Used [0-9]* registers') $(grep -A1 'k_trace_frameILb0' build/rtb_trace_$name.log | grep -o '[0-9]* bytes spill stores')"
done
