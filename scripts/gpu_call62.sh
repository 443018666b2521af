#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 8,64,8 -- "" "9=2" "9=8" 2>&1 | grep shape | cut -c1-42,76-130,150-200
