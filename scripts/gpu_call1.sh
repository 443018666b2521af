#!/bin/bash
# GPU call: parity tests (one process per file: a device-side trap poisons only that file), layer timings across
# shard shapes with PDL off/on, short bench runs.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
: > gpurun_out/t1.log
for f in tests/test_gpu_quest.py tests/test_gpu_attention.py tests/test_gpu_quest_manager.py tests/test_gpu_h2o_manager.py tests/test_gpu_tp2.py; do
  echo "=== $f" >> gpurun_out/t1.log
  timeout 900 python -m pytest "$f" -m gpu -q --timeout 600 -x 2>&1 | tail -40 >> gpurun_out/t1.log
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" >> gpurun_out/t1.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t1.log
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "7=1" "9=8" "7=1,9=8" "7=1,1=128" "5=1" "5=2" > gpurun_out/exp1.log 2>&1; echo "rc=$?" >> gpurun_out/exp1.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload cfg5 > gpurun_out/bench1.log 2>&1; echo "rc=$?" >> gpurun_out/bench1.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload cfg5 --pdl 1 > gpurun_out/bench1_pdl.log 2>&1; echo "rc=$?" >> gpurun_out/bench1_pdl.log
grep -E "^===|passed|failed|error|Error|smoke" gpurun_out/t1.log | head -40; cat gpurun_out/exp1.log; tail -c 1500 gpurun_out/bench1.log; tail -c 600 gpurun_out/bench1_pdl.log
