#!/bin/bash
cd "$(dirname "$0")/.."
for w in 4 2; do
  echo "== world $w default"; BAND_ROWS=4 python scripts/band_times.py $w 16777216 2>&1 | head -2
  echo "== world $w RTB200_TERMS=1"; BAND_ROWS=4 RTB200_TERMS=1 python scripts/band_times.py $w 16777216 2>&1 | head -2
  echo "== world $w one chunk (chunk 64 Mi)"; BAND_ROWS=4 python scripts/band_times.py $w 67108864 2>&1 | head -2
done
echo "== world 1 RTB200_TERMS=1"; BAND_ROWS=4 RTB200_TERMS=1 python scripts/band_times.py 1 16777216 2>&1 | head -1
echo "== world 1 default"; BAND_ROWS=4 python scripts/band_times.py 1 16777216 2>&1 | head -1
