#!/bin/bash
mkdir -p gpurun_out /tmp/ncu
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_datasets.py -m gpu -q -x 2>&1 | tail -4
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_m2.json 2> gpurun_out/bench_m2.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_m2.json')); print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'll', d['ll_after'], 'loglik GB/s', round(d['loglik_gbs'],1)); print({k['group']: k['ms_per_sweep'] for k in d['kernels']})"
timeout 600 ncu --set full --clock-control none -k "regex:k_item_ll" --launch-skip 3 --launch-count 1 -f -o /tmp/ncu/ll python tools/sweep_only.py 3 1 > gpurun_out/ncu_ll.log 2>&1
python tools/ncu_summary.py /tmp/ncu/ll.ncu-rep > gpurun_out/r02_ncu_item_ll_v3.txt; cat gpurun_out/r02_ncu_item_ll_v3.txt
