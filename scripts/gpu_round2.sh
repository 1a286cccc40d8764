#!/bin/bash
# One gpurun call of round 2: GPU tests, kernel A/B, bench lines.  usage: scripts/gpu_round2.sh [skip-tests]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt
if [ "$1" != "skip-tests" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  tail -15 gpurun_out/pytest_gpu.log
fi
timeout 600 python scripts/ab_kernels.py 3 > gpurun_out/ab_c3.txt 2>&1; cat gpurun_out/ab_c3.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_c3.json; tail -5 gpurun_out/bench_c3.err
