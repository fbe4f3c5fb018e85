#!/bin/bash
# Round 2, ncu visit: launch list of 2 sweeps at the bench's warm-up state (5 warm-up sweeps), then --set full captures. Plain runs first.
mkdir -p gpurun_out /tmp/ncu
timeout 300 python tools/sweep_only.py 5 2 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
cat gpurun_out/plain.log | tail -1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches.csv python tools/sweep_only.py 5 2 > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r02_launches.csv
cap() {  # name, kernel regex, skip, count, script...
  name=$1; rx=$2; skip=$3; cnt=$4; shift 4
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$rx" --launch-skip $skip --launch-count $cnt -f -o /tmp/ncu/r02_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"; tail -1 gpurun_out/ncu_$name.log
}
# 5 warm-up sweeps: skip the launches of the initial draw and of the first 5 sweeps (per sweep: 3 item ESS, 10 GEMM launches, 2 persistent Cholesky)
cap ess "k_ess_items" 15 3 python tools/sweep_only.py 5 2
cap chol "k_chol_persistent" 11 2 python tools/sweep_only.py 5 2
cap gemm "gemm_f64_kernel" 75 14 python tools/sweep_only.py 5 2
cap theta1 "k_ess_theta|k_theta_table|k_item_ll" 15 3 python tools/sweep_only.py 5 2
timeout 300 python tools/shards_only.py 5 1 > gpurun_out/plain_shards.log 2>&1 && tail -1 gpurun_out/plain_shards.log &&
cap thetaN "k_theta_partials|k_theta_decide|k_theta_table|k_pack_items" 40 10 python tools/shards_only.py 5 1
for r in ess chol gemm theta1 thetaN; do python tools/ncu_summary.py /tmp/ncu/r02_$r.ncu-rep > gpurun_out/r02_ncu_$r.txt 2> gpurun_out/ncu_summary_$r.err; wc -l gpurun_out/r02_ncu_$r.txt; done
python tools/summarize_launches.py gpurun_out/r02_launches.csv > gpurun_out/r02_launches_summary.txt 2>&1; tail -25 gpurun_out/r02_launches_summary.txt
ls -la /tmp/ncu | awk '{print $5, $9}'; du -sh gpurun_out
