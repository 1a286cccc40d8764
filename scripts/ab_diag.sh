#!/bin/bash
# Runs scripts/diag_queue.py (counters build: traversal statistics) once per variant library under variants/.
cd "$(dirname "$0")/.."
for lib in variants/librtb200_*.so; do echo $lib; RTB200_LIB=$PWD/$lib python scripts/diag_queue.py 2>&1 | grep -E "k2:|3840" | cut -c1-400; done
