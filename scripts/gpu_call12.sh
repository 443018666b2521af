#!/bin/bash
# GPU call 12: how many ring stages does the attention kernel need (a fused score->attention launch leaves it fewer)?
mkdir -p gpurun_out
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,8,1 -- "" "11=5" "11=4" "11=3" "11=2" "0=8" > gpurun_out/exp12.log 2>&1; cat gpurun_out/exp12.log | cut -c1-200
