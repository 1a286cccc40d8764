#!/bin/bash
# ncu launch list of the bench (after the same command exited 0 without ncu).  usage: gpu_launchlist.sh [name]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-r2}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${N}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python scripts/launch_summary.py gpurun_out/${N}_launches.csv > gpurun_out/${N}_launches.txt; head -40 gpurun_out/${N}_launches.txt
