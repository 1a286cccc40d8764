#!/bin/bash
# ncu launch list (time, active lanes, warp instructions) of the wavefront kernels of one 4K frame; for gpurun.
# usage: scripts/kernel_times.sh [nb_ray]
cd "$(dirname "$0")/.."
NR=${1:-4}
python scripts/profile_frame.py 3840 2160 $NR 3 > gpurun_out/plain2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum --clock-control none \
    -k regex:"k_primary|k_shadow|k_shade" -s 5 -c 5 --csv --log-file gpurun_out/l2.csv python scripts/profile_frame.py 3840 2160 $NR 3 > gpurun_out/ncu_l.log 2>&1
python - <<'PY'
import csv
rows = list(csv.reader(open('gpurun_out/l2.csv')))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
h = rows[hi]; out = {}
for r in rows[hi + 1:]:
    out.setdefault(r[h.index('Kernel Name')][:28], {})[r[h.index('Metric Name')].split('.')[0][-14:]] = r[h.index('Metric Value')]
for k, v in out.items(): print(k, v)
PY
