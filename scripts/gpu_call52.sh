#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 600 2>&1 | tail -2
timeout 900 python scripts/exp_layer.py 1,32,8,32768,2048 2,32,8,32768,2048 4,32,8,32768,2048 1,64,8,131072,4096 2,64,8,131072,4096 4,64,8 8,64,8 16,64,8 16,8,1 16,16,2 16,32,4 32,28,4 2>&1 | grep shape | cut -c1-42,76-130,150-200
