#!/bin/bash
# ncu --set full of the speculative rebuild's passes (count, fill, sort), config 3 and config 4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/e2e_breakdown.py 3 1 > gpurun_out/plain_e2e3.log 2>&1 || { tail -5 gpurun_out/plain_e2e3.log; exit 1; }
for c in 3 4; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_grid_(raster|sort)$' -s 6 -c 3 -f -o gpurun_out/r2_raster_c$c \
    python scripts/e2e_breakdown.py $c 1 > gpurun_out/ncu_raster$c.log 2>&1; echo "ncu raster $c rc=$?"
done
ls -la gpurun_out/*.ncu-rep
