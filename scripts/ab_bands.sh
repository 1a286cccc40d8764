#!/bin/bash
# per-rank (world 8) and full-frame kernel times for every variant library under variants/.  For gpurun.
cd "$(dirname "$0")/.."
for lib in variants/librtb200_*.so; do
  echo "== $lib"
  RTB200_LIB=$PWD/$lib python scripts/band_times.py 8 16777216 2>&1 | grep -E "rank [0145] " | cut -c1-140
  RTB200_LIB=$PWD/$lib python scripts/band_times.py 1 16777216 2>&1 | grep -E "rank 0 " | cut -c1-140
done
