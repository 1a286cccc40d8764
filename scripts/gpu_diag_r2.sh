#!/bin/bash
# Round-2 diagnostics in ONE gpurun call: full ncu captures (with source) of the three hot kernels of the 4K/16spp frame,
# the round statistics of k_shadow_grid (RTB_DIAG_GRID build), and compute-sanitizer memcheck / racecheck of a small run.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/diag_gpu.txt
python scripts/profile_frame.py 3840 2160 16 3 > gpurun_out/plain2.log 2>&1 || { tail -5 gpurun_out/plain2.log; exit 1; }
tail -1 gpurun_out/plain2.log
for spec in k_primary:13 k_shadow_grid:12 k_shade:12; do
  K=${spec%%:*}; S=${spec##*:}
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -f -o gpurun_out/r2a_$K \
      python scripts/profile_frame.py 3840 2160 16 3 > gpurun_out/ncu_$K.log 2>&1
  echo "ncu $K rc=$?"
done
if [ -f variants/librtb200_diag.so ]; then
  RTB200_LIB=$PWD/variants/librtb200_diag.so python scripts/diag_grid.py > gpurun_out/diag_grid.txt 2>&1; cat gpurun_out/diag_grid.txt
fi
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_run.py > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; tail -4 gpurun_out/sanitizer_$tool.log
done
