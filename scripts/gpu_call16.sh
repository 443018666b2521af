#!/bin/bash
# GPU call 16: two-wave one-launch layer: parity (layer + attention + managers), timings, timelines.
mkdir -p gpurun_out
: > gpurun_out/t16.log
for f in tests/test_gpu_quest_layer.py tests/test_gpu_attention.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t16.log
  timeout 600 python -m pytest "$f" -m gpu -x -q --timeout 200 2>&1 | tail -25 >> gpurun_out/t16.log
done
grep -E "^===|passed|failed|Error|error" gpurun_out/t16.log | head -20
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "12=1" > gpurun_out/exp16.log 2>&1; cat gpurun_out/exp16.log | cut -c1-230
SHAPE=16,64,8 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
SHAPE=16,8,1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
