#!/bin/bash
mkdir -p gpurun_out
for C in 3 3b 4 5; do
  timeout 900 python bench.py --config $C --steps 8 > gpurun_out/r2_bench_c${C}_1gpu.json 2> gpurun_out/r2_bench_c${C}_1gpu.err; echo "bench $C rc=$?"
done
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/ref.err; echo "reference arm rc=$?"; tail -c 600 gpurun_out/r2_bench_reference_arm.json
