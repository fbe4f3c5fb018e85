#!/bin/bash
# One `ncu --set full` capture of the three item ESS launches of a warm C3 sweep (after the same command has run clean without ncu).
tag=${1:-full}
mkdir -p gpurun_out
timeout 120 python tools/sweep_only.py 2 1 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
timeout 150 ncu --set full --clock-control none --import-source on -k "regex:k_ess_items" --launch-skip 6 --launch-count 3 -f -o gpurun_out/${tag}_ess \
    python tools/sweep_only.py 2 1 > gpurun_out/ncu_ess.log 2>&1; echo "ncu ess rc=$?"
ls -la gpurun_out/*.ncu-rep
