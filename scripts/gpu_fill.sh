#!/bin/bash
for b in 4 3 2; do
  echo "== shadow blocks per SM $b"
  RTB200_SHADOW_BLOCKS=$b python scripts/ab_kernels.py 3
  RTB200_SHADOW_BLOCKS=$b python scripts/ab_kernels.py 3b
done
