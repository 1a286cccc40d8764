#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
bash scripts/ab_variant_libs.sh 3 > gpurun_out/ab_variants.txt 2>&1; cat gpurun_out/ab_variants.txt
python scripts/profile_frame.py 3840 2160 16 3 > gpurun_out/plain2.log 2>&1 || { tail -5 gpurun_out/plain2.log; exit 1; }
for spec in k_shadow_grid2:12 k_primary2:13; do
  K=${spec%%:*}; S=${spec##*:}
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -f -o gpurun_out/r2b_$K \
      python scripts/profile_frame.py 3840 2160 16 3 > gpurun_out/ncu_$K.log 2>&1
  echo "ncu $K rc=$?"
done
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
