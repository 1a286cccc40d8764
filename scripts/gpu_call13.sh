#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AB_PAIRS="2,2 3,2"
( timeout 300 python scripts/ab_kernels.py 3; timeout 300 python scripts/ab_kernels.py 3b; timeout 300 python scripts/ab_kernels.py 4 ) 2>&1
timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
