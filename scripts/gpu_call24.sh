#!/bin/bash
# GPU call 24: evict-first L2 hints on the streaming copies: parity tests, launch-order experiment, layer timings with / without.
mkdir -p gpurun_out
: > gpurun_out/t24.log
for f in tests/test_gpu_quest_layer.py tests/test_gpu_attention.py tests/test_gpu_quest.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t24.log
  timeout 600 python -m pytest "$f" -m gpu -x -q --timeout 200 2>&1 | tail -25 >> gpurun_out/t24.log
done
grep -E "^===|passed|failed|Error|error" gpurun_out/t24.log | head -24
SHAPE=16,64,8 timeout 300 python scripts/exp_boundary.py 2>&1 | tail -5
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,8,1 1,32,8,32768,2048 -- "" "13=1" > gpurun_out/exp24.log 2>&1; cat gpurun_out/exp24.log | cut -c1-230
