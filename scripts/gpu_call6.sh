#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AB_PAIRS="2,2"
( python scripts/ab_kernels.py 3; RTB200_NO_COVER_ORDER=1 python scripts/ab_kernels.py 3 ) 2>&1 | grep -v identical > gpurun_out/ab_cover.txt; cat gpurun_out/ab_cover.txt
( BAND_ROWS=4 python scripts/band_times.py 8 16777216 8388608; BAND_ROWS=4 RTB200_TERMS=1 python scripts/band_times.py 8 8388608 ) 2>&1 | awk '{ov+=$9; n++; print} END {}' > gpurun_out/band8.txt; cat gpurun_out/band8.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "sequence or speculative" 2>&1 | tail -3
timeout 300 ./ray-tracer-rust_b200/rt_main assets/models/big_bunny.obj --orbit 120 --out /tmp/orb.png | tail -1
timeout 600 python bench.py --config 5 --steps 3 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; tail -c 900 gpurun_out/bench_c5.json
