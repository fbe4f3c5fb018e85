#!/bin/bash
# Round 2, visit K: batched chains with the divided CTA budget; chains tests; e2e.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_entry.py -q -x 2>&1 | tail -4
for b in 48 96 24; do
  GPIRT_CHAINS_CHOL_CTAS=$b timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_k_b$b.json 2> gpurun_out/bench_k_b$b.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_k_b$b.json')); print('budget $b', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['e2e']['calls'], 'chains', round(d['batched_chains']['value'],1))"
done
