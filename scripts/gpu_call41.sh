#!/bin/bash
mkdir -p gpurun_out
echo "## score+select timeline (SHAPE=16,8,1)"; SHAPE=16,8,1 timeout 200 python scripts/exp_score_times.py 2>&1 | tail -8
echo; echo "## one-launch layer timeline (SHAPE=16,8,1)"; SHAPE=16,8,1 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -32
