#!/bin/bash
mkdir -p gpurun_out
SHAPE=16,64,8 timeout 300 python scripts/exp_boundary.py 2>&1 | tail -6
SHAPE=16,8,1 timeout 300 python scripts/exp_boundary.py 2>&1 | tail -6
