#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/exp_layer.py 1,32,8,32768,2048 4,64,8 -- "" "4=1" "4=2" "4=4" "15=1" "15=4" "3=1" "3=2" "0=1" "0=2" 2>&1 | grep shape | cut -c1-110
