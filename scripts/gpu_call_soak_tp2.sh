#!/bin/bash
# Soak of the 2-rank parity tests (gpurun --gpus 2): both push protocols, one-launch push, several times.
N=${1:-3}
mkdir -p gpurun_out
for i in $(seq 1 $N); do
  timeout 600 python -m pytest tests/test_gpu_tp2.py -m gpu -q --timeout 500 -p no:cacheprovider 2>&1 | tail -1
done
