#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_quest_layer.py -m gpu -x -q --timeout 200 2>&1 | tail -2
timeout 600 python scripts/exp_layer.py 16,8,1 16,16,2 16,32,4 16,64,8 1,32,8,32768,2048 -- "" "14=500" "14=8000" > gpurun_out/exp30.log 2>&1; cat gpurun_out/exp30.log | cut -c1-230
SHAPE=16,8,1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -8
