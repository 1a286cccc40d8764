#!/bin/bash
# One gpurun call: ncu --set full of one kernel of the 4K/16spp frame (after the same command exited 0 without ncu).
# usage: scripts/gpu_profile.sh [kernel regex = k_shade] [launches to skip = 12] [count = 1] [report name = prof]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K=${1:-k_shade}; S=${2:-12}; C=${3:-1}; O=${4:-prof}
python scripts/profile_frame.py 3840 2160 16 3 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c $C -f -o gpurun_out/$O \
    python scripts/profile_frame.py 3840 2160 16 3 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
