#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/profile_frame.py 3840 2160 1 5 288 8 > gpurun_out/plain4.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_shadow_grid$' -s 4 -c 1 -f -o gpurun_out/r2_k_shadow_grid_c4 \
    python scripts/profile_frame.py 3840 2160 1 5 288 8 > gpurun_out/ncu_sg4.log 2>&1; echo "ncu 4 rc=$?"
