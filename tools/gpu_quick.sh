#!/bin/bash
# tests + bench line (no cpu baseline), summary printed
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -15
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ['value','ms_per_step','gpu_launches','ll_after','mean_evals','loglik_gbs']}, d['e2e']['value'])
for k in d['kernels']: print(k['group'], k['ms_per_sweep'], k['achieved'], k['unit'], k['frac'])
PY
