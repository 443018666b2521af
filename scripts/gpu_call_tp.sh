#!/bin/bash
# Multi-GPU call (gpurun --gpus N): 2-rank parity test, then bench at N ranks with the NVLink push epilogue (PDL off / on)
# and with the NCCL all-gather.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/tp_gpus.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_tp2.py -m gpu -q --timeout 500 > gpurun_out/tp2_test.log 2>&1; echo "tp2 test rc=$?" >> gpurun_out/tp2_test.log
tail -3 gpurun_out/tp2_test.log
for mode in push:0 push:1 nccl:0; do
  coll=${mode%%:*}; pdl=${mode##*:}
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 --collective $coll --pdl $pdl > gpurun_out/bench_n${N}_${coll}_pdl${pdl}.log 2>&1
  echo "bench N=$N $coll pdl=$pdl rc=$?"
  grep '^{' gpurun_out/bench_n${N}_${coll}_pdl${pdl}.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(' value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'score_us', round(d['roofline']['avg_launch_ms']*1e3,1), 'attn_us', round(d['roofline_attention']['avg_launch_ms']*1e3,1))
"
done
