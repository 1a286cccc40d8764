#!/bin/bash
# ncu --set full of the speculative rebuild's passes (count, fill, sort), config 3 and config 4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/e2e_breakdown.py 3 1 > gpurun_out/plain_e2e3.log 2>&1 || { tail -5 gpurun_out/plain_e2e3.log; exit 1; }
# launches that match before the steady new-table loop of e2e_breakdown.py (a table's third frame rebuilds finer, the first new table
# after that builds synchronously): 30 on config 3 (count, fill, sort per rebuild), 25 on config 4 (count, fill)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_grid_(raster|sort)$' -s 33 -c 3 -f -o gpurun_out/r2_raster_c3 \
    python scripts/e2e_breakdown.py 3 1 > gpurun_out/ncu_raster3.log 2>&1; echo "ncu raster 3 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_grid_(raster|sort)$' -s 27 -c 2 -f -o gpurun_out/r2_raster_c4 \
    python scripts/e2e_breakdown.py 4 1 > gpurun_out/ncu_raster4.log 2>&1; echo "ncu raster 4 rc=$?"
ls -la gpurun_out/*.ncu-rep
