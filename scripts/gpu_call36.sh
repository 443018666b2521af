#!/bin/bash
# GPU call 36: loader look-ahead by bytes + row tails in single units: parity, sweep, timeline.
mkdir -p gpurun_out
: > gpurun_out/t36.log
for f in tests/test_gpu_quest.py tests/test_gpu_quest_layer.py tests/test_gpu_quest_manager.py; do
  echo "=== $f" >> gpurun_out/t36.log
  timeout 600 python -m pytest "$f" -m gpu -x -q --timeout 200 2>&1 | tail -25 >> gpurun_out/t36.log
done
grep -E "^===|passed|failed|Error|error" gpurun_out/t36.log | head -12
timeout 900 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,8,1 1,32,8,32768,2048 -- "" "16=1" "15=4,16=1" "15=2" "16=10" "16=37" "4=2" "4=2,16=1" > gpurun_out/exp36.log 2>&1; cat gpurun_out/exp36.log | cut -c1-230
SHAPE=16,64,8 CLAIMS=1 timeout 200 python scripts/exp_score_times.py 2>&1 | tail -19 | head -8
