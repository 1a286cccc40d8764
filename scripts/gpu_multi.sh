#!/bin/bash
# Multi-GPU checks (gpurun --gpus N): the C-ABI multi-device tests, then bench.py under torchrun as the driver launches it.
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_$N.txt
timeout 900 python -m pytest tests/test_gpu_bands.py -m gpu -q -k "multi or peer" > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_multi_$N.log
for cfg in 3 5; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 5 --warmup 3 --config $cfg \
   > gpurun_out/bench_c${cfg}_${N}gpu.json 2> gpurun_out/bench_c${cfg}_${N}gpu.err; echo "bench config $cfg rc=$?"; tail -c 1500 gpurun_out/bench_c${cfg}_${N}gpu.json; tail -3 gpurun_out/bench_c${cfg}_${N}gpu.err
done
timeout 300 ./ray-tracer-rust_b200/rt_main assets/models/big_bunny.obj --gpus $N --width 3840 --height 2160 --spp 16 --out gpurun_out/rt_main_${N}gpu.png | tee gpurun_out/rt_main_${N}gpu.txt
timeout 300 ./ray-tracer-rust_b200/rt_main assets/models/big_bunny.obj --gpus $N --orbit 120 --out /tmp/orb.png | tee -a gpurun_out/rt_main_${N}gpu.txt
