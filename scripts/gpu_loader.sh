#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nproc; free -g | head -2; df -h /tmp | tail -1
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "parallel_import" 2>&1 | tail -3
timeout 600 python scripts/loader_bench.py 8 gpurun_out/r2_loader_k8.json 2>&1 | tail -2
timeout 1200 python scripts/loader_bench.py 64 gpurun_out/r2_loader_k64.json 2>&1 | tail -2
