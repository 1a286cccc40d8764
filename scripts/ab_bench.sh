#!/bin/bash
# Runs bench.py (no CPU baseline) once per variant library under variants/ and prints ms/frame.  For gpurun.
cd "$(dirname "$0")/.."
for lib in variants/librtb200_*.so; do
  RTB200_LIB=$PWD/$lib python bench.py --steps ${STEPS:-4} --warmup 3 --no-cpu-baseline 2>>gpurun_out/ab_err.log | tail -1 | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib', round(d['ms_per_step'],2), 'ms', round(d['value']), 'Mrays/s box/ray', round(d['traversal']['box_tests_per_ray'],2))"
done
