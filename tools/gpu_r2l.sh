#!/bin/bash
# Round 2, visit L: register budget of the persistent Cholesky (GPIRT_CHOL_MINB) x CTA budget, in the sweep and alone.
mkdir -p gpurun_out
for mb in 1 2 3; do echo "== MINB=$mb alone"; GPIRT_CHOL_MINB=$mb GPIRT_CHOL_CTAS=128 timeout 300 python tools/chol_check.py 2>&1 | grep "chol_K"; done
for mb in 1 2 3; do for c in 40 48 64; do
  GPIRT_CHOL_MINB=$mb GPIRT_CHOL_CTAS=$c timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_l_$mb_$c.json 2> gpurun_out/bench_l.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_l_$mb_$c.json')); print('minb $mb ctas $c', 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'll', d['ll_after'])"
done; done
