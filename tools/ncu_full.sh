#!/bin/bash
# One `ncu --set full` capture of the kernels that dominate a C3 sweep (after the same command has run clean without ncu).
# usage: tools/ncu_full.sh <tag>; writes gpurun_out/<tag>_*.ncu-rep
tag=${1:-full}
mkdir -p gpurun_out
timeout 300 python tools/sweep_only.py 2 1 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
cap() {  # name, kernel regex, skip, count
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$2" --launch-skip $3 --launch-count $4 -f -o gpurun_out/${tag}_$1 \
    python tools/sweep_only.py 2 1 > gpurun_out/ncu_$1.log 2>&1; echo "ncu $1 rc=$?"
}
cap ess k_ess_items 6 3
cap theta "k_ess_theta|k_theta_table|k_item_ll" 6 3
cap gemm "gemm_f64_kernel" 150 6
cap panel "k_chol_panel|k_trtri_diag" 40 4
ls -la gpurun_out/*.ncu-rep
