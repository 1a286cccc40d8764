#!/bin/bash
# end-to-end step of a band share with and without the light-grid rebuild, then the GPU parity tests
mkdir -p gpurun_out
{
  timeout 300 python scripts/e2e_breakdown.py 3 1 8
  timeout 300 python scripts/e2e_breakdown.py 4 1 8
  timeout 300 python scripts/e2e_breakdown.py 3b 1 8
} > gpurun_out/fill_ab.log 2>&1
cat gpurun_out/fill_ab.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
