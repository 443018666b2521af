#!/bin/bash
# GPU call: ncu launch list + one full capture of the two hot kernels (cfg5, eager launches, 2 physical layers).
# Each ncu run follows the same command exiting 0 without ncu in this call (B200_PROFILING.md).
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --phys-layers 2 --workload cfg5 ${BENCH_EXTRA}"
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"quest|sparse_decode|store_kv|h2o" -c 600 --csv \
    --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/r2_ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"quest_page_score|sparse_decode_attn" -s 40 -c 4 \
    -o gpurun_out/r2_prof $CMD > gpurun_out/r2_ncu_full.log 2>&1
echo "full capture rc=$?"
tail -3 gpurun_out/r2_ncu_plain.log | cut -c1-600
