#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/exp_layer.py 16,32,4 16,16,2 8,64,8 32,32,8,65536,4096 64,32,8,32768,2048 -- "" > gpurun_out/exp25.log 2>&1; cat gpurun_out/exp25.log | cut -c1-230
