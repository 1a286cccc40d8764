#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for lib in variants/librtb200_split*.so; do
  n=$(basename $lib .so)
  RTB200_LIB=$PWD/$lib ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_grid -c 60 --csv --log-file gpurun_out/$n.csv python scripts/profile_frame.py 1920 1080 16 2 > /dev/null 2>&1
  echo "== $n"; python scripts/launch_summary.py gpurun_out/$n.csv | head -8
done
