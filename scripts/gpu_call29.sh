#!/bin/bash
mkdir -p gpurun_out
SHAPE=16,8,1 NOWAVES=1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
SHAPE=16,8,1 CLAIMS=1 timeout 200 python scripts/exp_score_times.py 2>&1 | tail -6
