#!/bin/bash
# ncu --set full of the three hot kernels (config 3; k_shadow_grid also on 3b and 4) in the FIFTH frame of a table (lists ordered on the second, grids refined on the third), each after the same command exited 0 without ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/profile_frame.py 3840 2160 16 5 > gpurun_out/plain3.log 2>&1 || { tail -5 gpurun_out/plain3.log; exit 1; }
for spec in k_primary:37 k_shadow_grid:36 k_shade:36; do
  K=${spec%%:*}; S=${spec##*:}
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^$K\$" -s $S -c 1 -f -o gpurun_out/r2_$K \
      python scripts/profile_frame.py 3840 2160 16 5 > gpurun_out/ncu_$K.log 2>&1; echo "ncu $K rc=$?"
done
python scripts/profile_frame.py 3840 2160 16 5 576 > gpurun_out/plain3b.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_shadow_grid$' -s 36 -c 1 -f -o gpurun_out/r2_k_shadow_grid_c3b \
    python scripts/profile_frame.py 3840 2160 16 5 576 > gpurun_out/ncu_sg3b.log 2>&1; echo "ncu 3b rc=$?"
python scripts/profile_frame.py 3840 2160 1 5 288 8 > gpurun_out/plain4.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_shadow_grid$' -s 4 -c 1 -f -o gpurun_out/r2_k_shadow_grid_c4 \
    python scripts/profile_frame.py 3840 2160 1 5 288 8 > gpurun_out/ncu_sg4.log 2>&1; echo "ncu 4 rc=$?"
ls -la gpurun_out/*.ncu-rep
