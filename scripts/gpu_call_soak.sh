#!/bin/bash
# Soak: the GPU suite several times in a row (fresh process each), to catch flaky flag / ticket protocols.
N=${1:-5}
mkdir -p gpurun_out
for i in $(seq 1 $N); do
  timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 -p no:cacheprovider 2>&1 | tail -1
done
