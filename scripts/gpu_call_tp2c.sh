#!/bin/bash
# 2-GPU call: parity of both push protocols + the one-launch layer's push; bench with flag-in-data words and with flags.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tp2.py -m gpu -q --timeout 500 > gpurun_out/tp2_test.log 2>&1; echo "tp2 test rc=$?"; tail -15 gpurun_out/tp2_test.log | cut -c1-200
for extra in "" "--push-flags"; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --steps 20 --warmup 5 $extra > gpurun_out/bench_tp2c.log 2>&1
echo "bench N=2 [$extra] rc=$?"
grep '^{' gpurun_out/bench_tp2c.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
c=d.get('collective') or {}
print(' value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'score_us', round(d['roofline']['avg_launch_ms']*1e3,1), 'attn_us', round(d['roofline_attention']['avg_launch_ms']*1e3,1), 'coll_us', round(c.get('us_per_layer',0),1), 'other_us', round(c.get('other_us_per_layer',0),1))
" || tail -20 gpurun_out/bench_tp2c.log
done
