#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 600 2>&1 | tail -2
timeout 600 python scripts/exp_layer.py 1,32,8,32768,2048 4,32,8,32768,2048 16,32,8,65536,4096 16,64,8 16,8,1 2>&1 | grep shape | cut -c1-200
