#!/bin/bash
# GPU call 2: parity tests (one process per file), smoke, layer timings across shapes and knobs, timelines, full bench.
mkdir -p gpurun_out
: > gpurun_out/t2.log
for f in tests/test_gpu_quest.py tests/test_gpu_attention.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t2.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -60 >> gpurun_out/t2.log
done
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t2.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t2.log
timeout 900 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "9=8" "8=2" "5=1" "5=2" "1=128" > gpurun_out/exp2.log 2>&1; echo "rc=$?" >> gpurun_out/exp2.log
SHAPE=16,64,8 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp2_score_times.log 2>&1
SHAPE=16,8,1 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp2_score_times_tp8.log 2>&1
SHAPE=16,64,8 timeout 200 python scripts/exp_attn_times.py > gpurun_out/exp2_attn_times.log 2>&1
timeout 1500 python bench.py > gpurun_out/bench2.log 2> gpurun_out/bench2.err; echo "bench rc=$?" >> gpurun_out/bench2.err
grep -E "^===|passed|failed|smoke" gpurun_out/t2.log | head -30; cat gpurun_out/exp2.log | cut -c1-200; tail -c 600 gpurun_out/bench2.err
