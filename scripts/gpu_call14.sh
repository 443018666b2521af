#!/bin/bash
# GPU call 14: one-launch layer: parity against the two launches, then timings.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quest_layer.py -m gpu -x -q --timeout 200 > gpurun_out/t14.log 2>&1; echo "layer tests rc=$?"
tail -25 gpurun_out/t14.log | cut -c1-220
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" > gpurun_out/exp14.log 2>&1; cat gpurun_out/exp14.log | cut -c1-230
