#!/bin/bash
cd "$(dirname "$0")/.."
echo "== world 1, chunk sweep"; BAND_ROWS=4 python scripts/band_times.py 1 8388608 16777216 25165824 33554432 50331648 2>&1
echo "== world 4"; BAND_ROWS=4 python scripts/band_times.py 4 16777216 2>&1 | head -2
echo "== world 8"; BAND_ROWS=4 python scripts/band_times.py 8 16777216 2>&1 | head -2
