#!/bin/bash
mkdir -p gpurun_out
for t in "17=12" "17=16" "17=24" "17=30" "17=37"; do
  timeout 300 python bench.py --workload cfg2 --skip-layers 2 --no-cpu-baseline --tune "$t" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d.get('roofline_dense_layer',{}); print('cfg2 skip2 tune [$t] value', round(d['value'],1), 'dense layer us', round(r.get('avg_launch_ms',0)*1e3,1), 'frac', round(r.get('frac',0),3), 'attn us', round(d['roofline_attention']['avg_launch_ms']*1e3,1))"
done
