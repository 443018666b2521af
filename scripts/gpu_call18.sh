#!/bin/bash
# GPU call 18: single-wave one-launch layer with its fixes: all parity tests, timings, timeline at the TP-8 shape.
mkdir -p gpurun_out
: > gpurun_out/t18.log
for f in tests/test_gpu_quest_layer.py tests/test_gpu_attention.py tests/test_gpu_quest.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py tests/test_zz_no_alloc.py; do
  echo "=== $f" >> gpurun_out/t18.log
  timeout 600 python -m pytest "$f" -m gpu -x -q --timeout 200 2>&1 | tail -25 >> gpurun_out/t18.log
done
grep -E "^===|passed|failed|Error|error" gpurun_out/t18.log | head -24
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" > gpurun_out/exp18.log 2>&1; cat gpurun_out/exp18.log | cut -c1-230
SHAPE=16,8,1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
