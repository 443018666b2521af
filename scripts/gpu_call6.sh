#!/bin/bash
# GPU call 6: claim-size sweep of the score kernel at cfg5 / cfg3 shapes.
mkdir -p gpurun_out
timeout 900 python scripts/exp_layer.py 16,64,8 32,28,4 -- "" "4=2" "4=2,2=1" "4=3,2=1" "4=2,2=3" "3=8" "3=8,2=1" "3=2" "3=2,2=1" "4=2,3=8,2=1" > gpurun_out/exp6.log 2>&1; echo "rc=$?" >> gpurun_out/exp6.log
cat gpurun_out/exp6.log | cut -c1-200
