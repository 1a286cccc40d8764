#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
{
for v in "" nodepth; do
  if [ -n "$v" ]; then export RTB200_LIB=$PWD/variants/librtb200_$v.so; else unset RTB200_LIB; fi
  echo "== variant ${v:-head(depth)}"
  timeout 300 python scripts/e2e_breakdown.py 3 1 | grep -v pipelined
  timeout 300 python scripts/e2e_breakdown.py 3b 1 | grep -v pipelined
  timeout 300 python scripts/e2e_breakdown.py 4 1 | grep -v pipelined
done
} > gpurun_out/fill_ab.log 2>&1
cat gpurun_out/fill_ab.log
RTB200_LIB=$PWD/variants/librtb200_diag2.so python scripts/diag_grid2.py 2>&1 | tail -4
