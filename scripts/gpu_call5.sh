#!/bin/bash
# GPU call 5: even static cut / chunk sizes / 2-bit search rounds / attention split penalty: parity tests, sweep, timelines.
mkdir -p gpurun_out
: > gpurun_out/t5.log
for f in tests/test_gpu_quest.py tests/test_gpu_attention.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t5.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -60 >> gpurun_out/t5.log
done
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t5.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t5.log
timeout 900 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "10=1" "10=5" "10=13" "10=17" "3=4" "4=1" > gpurun_out/exp5.log 2>&1; echo "rc=$?" >> gpurun_out/exp5.log
SHAPE=16,64,8 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp5_score_times.log 2>&1
SHAPE=16,8,1 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp5_score_times_tp8.log 2>&1
SHAPE=16,64,8 timeout 200 python scripts/exp_attn_times.py > gpurun_out/exp5_attn_times.log 2>&1
grep -E "^===|passed|failed|smoke" gpurun_out/t5.log | head -30; cat gpurun_out/exp5.log | cut -c1-200
