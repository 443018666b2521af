#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quest.py tests/test_gpu_quest_layer.py tests/test_gpu_quest_manager.py -m gpu -x -q --timeout 200 2>&1 | tail -2
timeout 600 python scripts/exp_layer.py 16,64,8 32,28,4 16,32,4 16,8,1 1,32,8,32768,2048 -- "" > gpurun_out/exp39.log 2>&1; cat gpurun_out/exp39.log | cut -c1-230
