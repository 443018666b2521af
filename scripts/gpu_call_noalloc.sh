#!/bin/bash
# GPU call: the no-allocation entry point's test alone in a fresh process, then after every other GPU test.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"
tail -3 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/noalloc_all.log 2>&1; echo "all rc=$?"
tail -4 gpurun_out/noalloc_all.log
