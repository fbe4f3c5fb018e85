#!/bin/bash
# Round 2, final visit (1 GPU): what the driver runs at round end — GPU tests, smoke, both bench arms.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/f_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/f_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/f_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/f_smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/f_ref.log 2>&1; echo "ref rc=$?"; grep '"impl"' gpurun_out/f_ref.log | tail -1 | cut -c1-240
timeout 600 python bench.py > gpurun_out/f_bench.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/f_bench.log | tail -1 | cut -c1-240
