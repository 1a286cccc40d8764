#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AB_PAIRS="2,2"
for t in 0 1 2 3; do echo "TUNE=$t"; RTB200_TUNE=$t python scripts/ab_kernels.py 3 2>&1 | grep -v identical; done > gpurun_out/ab_tune.txt; cat gpurun_out/ab_tune.txt
RTB200_TUNE=3 python scripts/ab_kernels.py 4 2>&1 | grep -v identical
RTB200_TUNE=3 timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
