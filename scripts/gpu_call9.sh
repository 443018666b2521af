#!/bin/bash
# GPU call 9: designated incremental merger of the attention kernel: parity tests, sweep, timeline.
mkdir -p gpurun_out
: > gpurun_out/t9.log
for f in tests/test_gpu_attention.py tests/test_gpu_quest.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t9.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -60 >> gpurun_out/t9.log
done
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t9.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t9.log
timeout 900 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "10=1" "10=13" "10=17" > gpurun_out/exp9.log 2>&1; echo "rc=$?" >> gpurun_out/exp9.log
SHAPE=16,64,8 timeout 200 python scripts/exp_attn_times.py > gpurun_out/exp9_attn_times.log 2>&1
grep -E "^===|passed|failed|smoke" gpurun_out/t9.log | head -30; cat gpurun_out/exp9.log | cut -c1-200; head -12 gpurun_out/exp9_attn_times.log
