#!/bin/bash
# One gpurun --gpus 8 call: the bench line at N = 2, 4, 8 (torchrun, as the driver launches it).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for N in ${@:-2 4 8}; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
      bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err; echo "N=$N rc=$?"
  python - $N <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/scale_%s.json" % n).read().strip().splitlines()[-1])
    print("N=%s ms/frame %.3f e2e %.3f kernel_ms(rank0) %.3f" % (n, d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms"]))
except Exception as e:
    print("N=%s: no line (%s)" % (n, e))
PY
done
