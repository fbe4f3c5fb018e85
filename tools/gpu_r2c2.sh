#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/e2e_fixed_cost.py 2>&1 | tail -8 | tee gpurun_out/c2_e2e.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/c2_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/c2_tests.log
