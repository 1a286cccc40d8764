#!/bin/bash
cd "$(dirname "$0")/.."
export AB_PAIRS="2,2"
for f in 0.18 0.125 0.09 0.0625; do timeout 300 python scripts/ab_kernels.py 3 grid_cell_factor=$f 2>&1 | grep -v identical; done
for f in 0.25 0.125; do timeout 300 python scripts/ab_kernels.py 3b grid_cell_factor=$f 2>&1 | grep -v identical; done
for f in 0.25 0.125 0.0625; do timeout 300 python scripts/ab_kernels.py 4 grid_cell_factor=$f 2>&1 | grep -v identical; done
