#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/e2e_fixed_cost.py 2>&1 | tail -8 | tee gpurun_out/d2_e2e.log
timeout 900 python -m pytest tests/test_gpu_entry.py tests/test_gpu_shards_local.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/d2_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/d2_tests.log
