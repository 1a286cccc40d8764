#!/bin/bash
# config 4 on 8 GPUs (after the entry budget of refined grids was raised)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29788 bench.py --gpus 8 --steps 8 --warmup 3 --config 4 \
   > gpurun_out/r2_bench_c4_8gpu.json 2> gpurun_out/r2_bench_c4_8gpu.err; echo "bench config 4 x8 rc=$?"
python - <<'PY'
import json
j=json.loads(open("gpurun_out/r2_bench_c4_8gpu.json").read().strip().splitlines()[-1])
print("value %.0f ms %.3f e2e %.0f e2e_ms %.3f frac %.3f" % (j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e"]["ms_per_step"], j["roofline"]["frac"]), {k: j[k] for k in j if "match" in k})
PY
