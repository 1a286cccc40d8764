#!/bin/bash
bash scripts/gpu_ncu_raster.sh
