#!/bin/bash
# config 3 on 8 GPUs (after k_primary's split for single-chunk renders)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29738 bench.py --gpus 8 --steps 8 --warmup 3 --config 3 --no-cpu-baseline \
   > gpurun_out/r2_bench_c3_8gpu.json 2> gpurun_out/r2_bench_c3_8gpu.err; echo "bench config 3 x8 rc=$?"
python - <<'PY'
import json
j=json.loads(open("gpurun_out/r2_bench_c3_8gpu.json").read().strip().splitlines()[-1])
print("value %.0f ms %.3f e2e %.0f e2e_ms %.3f frac %.3f" % (j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e"]["ms_per_step"], j["roofline"]["frac"]), {k: round(v["ms"],3) for k,v in j["roofline"]["per_kernel"].items()}, {k: j[k] for k in j if "match" in k})
PY
