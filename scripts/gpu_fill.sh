#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
{
  python scripts/e2e_breakdown.py 3 1 8
  python scripts/e2e_breakdown.py 3b 1 | grep -v pipelined
  python scripts/e2e_breakdown.py 4 1 | grep -v pipelined
} > gpurun_out/fill_ab.log 2>&1
cat gpurun_out/fill_ab.log
