#!/bin/bash
mkdir -p gpurun_out
echo "## cfg5 score+select timeline (SHAPE=16,64,8)"; SHAPE=16,64,8 CLAIMS=1 timeout 200 python scripts/exp_score_times.py 2>&1 | tail -20
echo; echo "## cfg5 attention timeline"; SHAPE=16,64,8 timeout 200 python scripts/exp_attn_times.py 2>&1 | tail -14
echo; echo "## cfg5 one-launch layer timeline"; SHAPE=16,64,8 timeout 200 python scripts/exp_fused_times.py 2>&1 | tail -14
