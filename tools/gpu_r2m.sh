#!/bin/bash
# Round 2, visit M: the table-based Phi in k_item_ll / k_theta_table: accuracy, parity suite, bench.
mkdir -p gpurun_out
timeout 1300 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_m.json 2> gpurun_out/bench_m.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_m.json')); print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'll', d['ll_after'], 'loglik GB/s', round(d['loglik_gbs'],1)); print({k['group']: k['ms_per_sweep'] for k in d['kernels']})"
