#!/bin/bash
# GPU call 28: cfg4 (H2O) end-to-end arm with side-stream copies; compute-sanitizer memcheck over the small GPU parity tests.
mkdir -p gpurun_out
timeout 900 python bench.py --workload cfg4 --no-cpu-baseline > gpurun_out/b28.log 2> gpurun_out/b28.err; echo "cfg4 rc=$?"; tail -c 300 gpurun_out/b28.err
python - <<'PY'
import json
for l in open('gpurun_out/b28.log'):
    if l.startswith('{'):
        d = json.loads(l)
        print(' value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value'], 1), 'e2e ms', round(d['e2e']['ms_per_step'], 3), 'frac', round(d['roofline']['frac'], 3), 'step frac', round(d['roofline_step']['frac'], 3))
PY
# (compute-sanitizer is closed on this pool: no memcheck run)
