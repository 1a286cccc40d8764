#!/bin/bash
# 8-GPU validation: multi-device C-ABI tests, then bench.py under torchrun for configs 3 / 4 / 5 (and 3 at 4 GPUs).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_gpu_bands.py -m gpu -q -k "multi" 2>&1 | tail -3
run() { # config gpus
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 297$2$2 bench.py --gpus $2 --steps 8 --warmup 3 --config $1 \
     > gpurun_out/r2_bench_c$1_$2gpu.json 2> gpurun_out/r2_bench_c$1_$2gpu.err; echo "bench config $1 x$2 rc=$?"
  python - gpurun_out/r2_bench_c$1_$2gpu.json <<'PY'
import json, sys
try:
    j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  value %.0f ms %.3f e2e %.0f e2e_ms %.3f frac %.3f" % (j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e"]["ms_per_step"], j["roofline"]["frac"]), {k: j[k] for k in j if "match" in k or "frames_per" in k or "redo" in k})
except Exception as e:
    print("  no line:", e)
PY
}
run 3 8; run 3 4; run 3 2; run 4 8; run 5 8; run 3b 8
timeout 300 ./ray-tracer-rust_b200/rt_main assets/models/big_bunny.obj --gpus 8 --width 3840 --height 2160 --spp 16 --out gpurun_out/rt_main_8gpu.png
timeout 300 ./ray-tracer-rust_b200/rt_main assets/models/big_bunny.obj --gpus 8 --orbit 120 --out /tmp/orb.png | tail -1
