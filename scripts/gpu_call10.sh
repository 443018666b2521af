#!/bin/bash
# GPU call 10: all GPU parity tests on the final kernels, then the ncu launch list of a short eager bench run.
mkdir -p gpurun_out
: > gpurun_out/t10.log
for f in tests/test_gpu_attention.py tests/test_gpu_quest.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py; do
  echo "=== $f" >> gpurun_out/t10.log
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 2>&1 | tail -30 >> gpurun_out/t10.log
done
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t10.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t10.log
grep -E "^===|passed|failed|smoke" gpurun_out/t10.log | head -30
timeout 300 python scripts/exp_layer.py 16,64,8 16,8,1 -- "" > gpurun_out/exp10.log 2>&1; cat gpurun_out/exp10.log | cut -c1-200
bash scripts/gpu_call_ncu_list.sh
