#!/bin/bash
# N ranks: the one-launch layer forced on against the auto rule (two launches at N >= 4), same box, back to back
N=${1:-4}
mkdir -p gpurun_out
for m in on auto; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 \
  bench.py --gpus $N --steps 20 --warmup 5 --one-launch $m 2>/dev/null | grep '^{' | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('N=$N one-launch $m: value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'launches/layer', d['launches_per_sparse_layer'])"
done
