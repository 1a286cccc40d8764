#!/bin/bash
# config 3 on 2 GPUs at HEAD (sanity of the torchrun path after late changes), plus the multi-GPU C-ABI tests
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bands.py -m gpu -q 2>&1 | tail -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29722 bench.py --gpus 2 --steps 8 --warmup 3 --no-cpu-baseline \
   > gpurun_out/n2.json 2> gpurun_out/n2.err; echo "bench x2 rc=$?"
python - <<'PY'
import json
j=json.loads(open("gpurun_out/n2.json").read().strip().splitlines()[-1])
print("value %.0f ms %.3f e2e_ms %.3f frac %.3f" % (j["value"], j["ms_per_step"], j["e2e"]["ms_per_step"], j["roofline"]["frac"]), {k: j[k] for k in j if "match" in k})
PY
