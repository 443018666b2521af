#!/bin/bash
# GPU call 11: cheaper cut arithmetic of the attention kernel: parity tests + shape sweep.
mkdir -p gpurun_out
: > gpurun_out/t11.log
for f in tests/test_gpu_attention.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t11.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -30 >> gpurun_out/t11.log
done
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t11.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t11.log
grep -E "^===|passed|failed|smoke" gpurun_out/t11.log | head -30
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" > gpurun_out/exp11.log 2>&1; cat gpurun_out/exp11.log | cut -c1-200
