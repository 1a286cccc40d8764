#!/bin/bash
mkdir -p gpurun_out
{
for v in "" kl64 kl128 kl160; do
  if [ -n "$v" ]; then export RTB200_LIB=$PWD/variants/librtb200_$v.so; else unset RTB200_LIB; fi
  echo "== variant ${v:-head(96)}"
  timeout 300 python scripts/e2e_breakdown.py 4 1 | grep -v pipelined
  timeout 300 python scripts/e2e_breakdown.py 3 1 | grep -v pipelined
done
} > gpurun_out/fill_ab.log 2>&1
cat gpurun_out/fill_ab.log
