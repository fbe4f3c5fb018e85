#!/bin/bash
# Round 2, last ncu visit: launch list of 2 warm sweeps and --set full of the three item ESS kernels, final code state. Plain run first.
mkdir -p gpurun_out /tmp/ncu
timeout 300 python tools/sweep_only.py 5 2 > gpurun_out/n2_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/n2_plain.log; exit 1; }
tail -1 gpurun_out/n2_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_final.csv python tools/sweep_only.py 5 2 > gpurun_out/n2_list.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r02_launches_final.csv
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_ess_items" --launch-skip 15 --launch-count 3 -f -o /tmp/ncu/r02_ess_packed python tools/sweep_only.py 5 2 > gpurun_out/n2_ess.log 2>&1
echo "ncu ess rc=$?"; tail -1 gpurun_out/n2_ess.log
python tools/ncu_summary.py /tmp/ncu/r02_ess_packed.ncu-rep > gpurun_out/r02_ncu_ess_packed.txt 2> gpurun_out/n2_summary.err; wc -l gpurun_out/r02_ncu_ess_packed.txt
python tools/summarize_launches.py gpurun_out/r02_launches_final.csv > gpurun_out/r02_launches_final_summary.txt 2>&1; head -30 gpurun_out/r02_launches_final_summary.txt
du -sh gpurun_out
