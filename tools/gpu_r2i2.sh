#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do timeout 200 python tools/band_sweep.py 2>&1 | tail -1; done | cut -c1-140 | tee gpurun_out/i2_tau2.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/i2_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/i2_tests.log
timeout 600 python bench.py > gpurun_out/i2_bench.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/i2_bench.log | tail -1 | cut -c1-200
