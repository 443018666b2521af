#!/bin/bash
# GPU call: ncu launch list (gpu__time_duration) of a short eager bench run, after the same command exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --phys-layers 2 --workload cfg5"
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"quest|sparse_decode|store_kv|h2o" -c 1200 --csv \
    --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
echo "launch list rc=$?"
tail -2 gpurun_out/r2_ncu_plain.log | cut -c1-400
