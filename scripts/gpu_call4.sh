#!/bin/bash
# GPU call 4: lane-parallel score loader / look-ahead refill order: parity tests, shape sweep, timelines.
mkdir -p gpurun_out
: > gpurun_out/t4.log
for f in tests/test_gpu_quest.py tests/test_gpu_attention.py tests/test_gpu_quest_manager.py; do
  echo "=== $f" >> gpurun_out/t4.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -60 >> gpurun_out/t4.log
done
timeout 900 python scripts/exp_layer.py 16,64,8 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "4=1" "4=2" "4=8" "5=1" > gpurun_out/exp4.log 2>&1; echo "rc=$?" >> gpurun_out/exp4.log
SHAPE=16,64,8 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp4_score_times.log 2>&1
SHAPE=16,8,1 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp4_score_times_tp8.log 2>&1
SHAPE=16,8,1 timeout 200 python scripts/exp_attn_times.py > gpurun_out/exp4_attn_times_tp8.log 2>&1
grep -E "^===|passed|failed|smoke" gpurun_out/t4.log | head -30; cat gpurun_out/exp4.log | cut -c1-200
