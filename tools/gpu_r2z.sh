#!/bin/bash
# Round 2, visit Z (1 GPU): W on a second side stream beside K_post + chol(K_post) (on / off), timeline, then the full GPU suite.
mkdir -p gpurun_out
for f in 1 0 1 0; do GPIRT_FORK_W=$f timeout 200 python tools/band_sweep.py 2>&1 | tail -1 | sed "s/^/fork_w=$f /"; done | tee gpurun_out/z_fork.log
timeout 200 python tools/timeline_c3.py 2>&1 | tail -6 | tee gpurun_out/z_timeline.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/z_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/z_tests.log
