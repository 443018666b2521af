#!/bin/bash
# GPU call: one ncu --set full capture of the two hot kernels at one shape ($SHAPE), after the same command ran clean.
mkdir -p gpurun_out
TAG=${TAG:-shape}
python scripts/exp_ncu_shape.py > gpurun_out/ncu_${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"quest_page_score|sparse_decode_attn" -s 9 -c 6 \
    -o gpurun_out/r2_${TAG} python scripts/exp_ncu_shape.py > gpurun_out/ncu_${TAG}.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_${TAG}.log
