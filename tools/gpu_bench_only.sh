#!/bin/bash
# bench line only (no tests, no cpu baseline), per-group summary printed
mkdir -p gpurun_out
timeout 600 python bench.py --steps ${1:-10} --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ['value','ms_per_step','ll_after']}, 'e2e', d['e2e']['value'])
print(' '.join(f"{k['group']}={k['ms_per_sweep']:.3f}" for k in d['kernels']))
PY
