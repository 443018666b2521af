#!/bin/bash
# GPU call: parity tests, layer timings across shard shapes with PDL off/on, short bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t1.log
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,16,2 16,8,1 1,32,8,32768,2048 -- "" "7=1" > gpurun_out/exp1.log 2>&1; echo "rc=$?" >> gpurun_out/exp1.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload cfg5 > gpurun_out/bench1.log 2>&1; echo "rc=$?" >> gpurun_out/bench1.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload cfg5 --pdl 1 > gpurun_out/bench1_pdl.log 2>&1; echo "rc=$?" >> gpurun_out/bench1_pdl.log
tail -5 gpurun_out/t1.log; cat gpurun_out/exp1.log; tail -c 1500 gpurun_out/bench1.log; tail -c 600 gpurun_out/bench1_pdl.log
