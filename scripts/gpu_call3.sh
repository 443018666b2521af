#!/bin/bash
# GPU call 3: parity tests after the select / scheduler / attention-epilogue changes, shape sweep, timelines, cfg4.
mkdir -p gpurun_out
: > gpurun_out/t3.log
for f in tests/test_gpu_quest.py tests/test_gpu_attention.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t3.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -60 >> gpurun_out/t3.log
done
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t3.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t3.log
timeout 900 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "3=4" > gpurun_out/exp3.log 2>&1; echo "rc=$?" >> gpurun_out/exp3.log
SHAPE=16,64,8 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp3_score_times.log 2>&1
SHAPE=16,8,1 CLAIMS=1 timeout 200 python scripts/exp_score_times.py > gpurun_out/exp3_score_times_tp8.log 2>&1
SHAPE=16,64,8 timeout 200 python scripts/exp_attn_times.py > gpurun_out/exp3_attn_times.log 2>&1
SHAPE=16,8,1 timeout 200 python scripts/exp_attn_times.py > gpurun_out/exp3_attn_times_tp8.log 2>&1
timeout 600 python bench.py --workload cfg4 --no-cpu-baseline > gpurun_out/bench3_cfg4.log 2> gpurun_out/bench3_cfg4.err; echo "bench rc=$?" >> gpurun_out/bench3_cfg4.err
grep -E "^===|passed|failed|smoke" gpurun_out/t3.log | head -30; cat gpurun_out/exp3.log | cut -c1-200; tail -c 400 gpurun_out/bench3_cfg4.err
