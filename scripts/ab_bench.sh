#!/bin/bash
# Runs scripts/time_frame.py once per variant library under variants/.  For gpurun.
cd "$(dirname "$0")/.."
for lib in variants/librtb200_*.so; do
  RTB200_LIB=$PWD/$lib python scripts/time_frame.py $1 2>&1 | tail -4
done
