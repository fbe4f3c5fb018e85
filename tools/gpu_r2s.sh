#!/bin/bash
# Round 2, visit S (1 GPU): packed fp32x2 screening in the ESS item kernels — full GPU suite, bench.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/s_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/s_tests.log
timeout 600 python bench.py > gpurun_out/s_bench.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/s_bench.log | tail -1 | cut -c1-300
python - <<'PY'
import json
for line in open('gpurun_out/s_bench.log'):
    if line.startswith('{"metric"'):
        d = json.loads(line)
        print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'])
        print({x['group']: x['ms_per_sweep'] for x in d['kernels']})
PY
