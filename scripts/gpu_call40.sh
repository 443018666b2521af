#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/exp_layer.py 16,8,1 16,16,2 16,32,4 1,32,8,32768,2048 16,64,8 -- "" "1=128" "1=192" "1=384" "1=512" "10=1" "10=5" "10=13" > gpurun_out/exp40.log 2>&1; cat gpurun_out/exp40.log | cut -c1-150
