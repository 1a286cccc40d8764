#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AB_PAIRS="2,2"
for f in 0.35 0.25 0.18 0.125 0.09; do
  timeout 300 python scripts/ab_kernels.py 3 grid_cell_factor=$f 2>&1 | grep -v identical
done > gpurun_out/ab_cell.txt
cat gpurun_out/ab_cell.txt
