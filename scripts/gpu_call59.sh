#!/bin/bash
# N=2: the one-launch layer (auto) against two launches, same box, back to back
mkdir -p gpurun_out
for m in auto off auto off; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
  bench.py --gpus 2 --steps 20 --warmup 5 --one-launch $m 2>/dev/null | grep '^{' | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('one-launch $m: value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'launches/layer', d['launches_per_sparse_layer'])"
done
