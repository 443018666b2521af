#!/bin/bash
# GPU call 15: one-launch layer: parity, timings, per-CTA timeline.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quest_layer.py -m gpu -x -q --timeout 200 > gpurun_out/t15.log 2>&1; echo "layer tests rc=$?"
tail -3 gpurun_out/t15.log | cut -c1-220
timeout 600 python scripts/exp_layer.py 16,64,8 16,32,4 16,8,1 1,32,8,32768,2048 -- "" > gpurun_out/exp15.log 2>&1; cat gpurun_out/exp15.log | cut -c1-230
SHAPE=16,64,8 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
SHAPE=16,8,1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
