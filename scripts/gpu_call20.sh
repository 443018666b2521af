#!/bin/bash
# GPU call 20: sync-free host path of the Quest decode step: manager tests, e2e of cfg5 / cfg3 / cfg2.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_quest_manager.py tests/test_gpu_quest.py -m gpu -x -q --timeout 200 2>&1 | tail -4
for args in "--workload cfg5" "--workload cfg3" "--workload cfg2"; do
  timeout 600 python bench.py $args --no-cpu-baseline > gpurun_out/b20.log 2> gpurun_out/b20.err; echo "[$args] rc=$?"; tail -c 300 gpurun_out/b20.err
  python - <<'PY'
import json
for l in open('gpurun_out/b20.log'):
    if l.startswith('{'):
        d = json.loads(l)
        print(' value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value'], 1), 'e2e ms', round(d['e2e']['ms_per_step'], 3))
PY
done
