#!/bin/bash
# GPU call 26: the whole GPU suite in one process (as the round's end runs it), PDL re-test, ncu launch list.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/t26.log 2>&1; echo "pytest -m gpu rc=$?"; tail -3 gpurun_out/t26.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python scripts/exp_layer.py 16,64,8 16,8,1 -- "" "7=1" > gpurun_out/exp26.log 2>&1; cat gpurun_out/exp26.log | cut -c1-230
bash scripts/gpu_call_ncu_list.sh
