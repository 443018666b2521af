#!/bin/bash
# GPU call 17: one-launch layer with late selectors / blocked consumers: parity, timings, timelines.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_quest_layer.py -m gpu -x -q --timeout 200 2>&1 | tail -5
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "12=1" > gpurun_out/exp17.log 2>&1; cat gpurun_out/exp17.log | cut -c1-230
SHAPE=16,64,8 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
SHAPE=16,8,1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
echo "no waves:"
SHAPE=16,8,1 NOWAVES=1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
