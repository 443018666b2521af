#!/bin/bash
# GPU call 8: guarded-subtraction select search, single-head consumer path: parity tests + sweep + select timeline.
mkdir -p gpurun_out
: > gpurun_out/t8.log
for f in tests/test_gpu_quest.py tests/test_gpu_attention.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t8.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -60 >> gpurun_out/t8.log
done
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t8.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t8.log
timeout 900 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "5=1" > gpurun_out/exp8.log 2>&1; echo "rc=$?" >> gpurun_out/exp8.log
SHAPE=16,8,1 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp8_score_times_tp8.log 2>&1
grep -E "^===|passed|failed|smoke" gpurun_out/t8.log | head -30; cat gpurun_out/exp8.log | cut -c1-200; tail -4 gpurun_out/exp8_score_times_tp8.log
