#!/bin/bash
mkdir -p gpurun_out
for c in 3 4; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_e2e$c.csv python scripts/e2e_breakdown.py $c 1 > gpurun_out/ncu_e2e$c.log 2>&1; echo rc=$?
done
