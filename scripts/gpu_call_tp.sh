#!/bin/bash
# Multi-GPU call (gpurun --gpus N): 2-rank parity test (when N >= 2), then bench at N ranks with the NVLink push
# epilogue and with the NCCL all-gather, PDL as given in $PDL (default 1).
N=${1:-2}
PDL=${PDL:-1}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/tp_gpus.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_tp2.py -m gpu -q --timeout 500 > gpurun_out/tp2_test.log 2>&1; echo "tp2 test rc=$?" >> gpurun_out/tp2_test.log
tail -5 gpurun_out/tp2_test.log
for coll in push nccl; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 --collective $coll --pdl $PDL > gpurun_out/bench_n${N}_${coll}.log 2>&1
  echo "bench N=$N $coll rc=$?"
  grep '^{' gpurun_out/bench_n${N}_${coll}.log | tail -1 | cut -c1-900
done
