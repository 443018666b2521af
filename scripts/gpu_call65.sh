#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/exp_layer.py 1,32,8,32768,2048 4,32,8,32768,2048 -- "" "1=128" "1=192" "17=6" "17=12" "17=7,1=128" 2>&1 | grep shape | cut -c1-42,76-130
