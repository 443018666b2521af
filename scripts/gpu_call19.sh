#!/bin/bash
# GPU call 19: bench plumbing of the one-launch layer (cfg3 auto = on; cfg5 auto = off, forced on for comparison).
mkdir -p gpurun_out
for args in "--workload cfg3" "--workload cfg3 --one-launch off" "--workload cfg5 --one-launch on" "--workload cfg2 --one-launch on"; do
  timeout 600 python bench.py $args --no-cpu-baseline --steps 10 > gpurun_out/b19.log 2> gpurun_out/b19.err; echo "[$args] rc=$?"; tail -c 300 gpurun_out/b19.err
  python - <<'PY'
import json
for l in open('gpurun_out/b19.log'):
    if l.startswith('{'):
        d = json.loads(l)
        print(' value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value'], 1), 'launches', d['gpu_launches'], d.get('launches_per_sparse_layer'),
              'score_us', round(d['roofline']['avg_launch_ms'] * 1e3, 1), 'attn_us', round(d['roofline_attention']['avg_launch_ms'] * 1e3, 1),
              'one', {k: (round(v, 4) if isinstance(v, float) else v) for k, v in (d.get('roofline_one_launch_layer') or {}).items() if k in ('avg_launch_ms', 'frac', 'two_launches_ms')})
PY
done
