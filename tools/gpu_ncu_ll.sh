#!/bin/bash
mkdir -p gpurun_out /tmp/ncu
timeout 300 python tools/sweep_only.py 3 1 > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_item_ll" --launch-skip 3 --launch-count 1 -f -o /tmp/ncu/ll python tools/sweep_only.py 3 1 > gpurun_out/ncu_ll.log 2>&1; echo "rc=$?"
python tools/ncu_summary.py /tmp/ncu/ll.ncu-rep > gpurun_out/r02_ncu_item_ll_v2.txt; cat gpurun_out/r02_ncu_item_ll_v2.txt
ncu -i /tmp/ncu/ll.ncu-rep --page details --section WarpStateStats --section SchedulerStats --section LaunchStats --section Occupancy 2>/dev/null | grep -E "Theoretical|Achieved|Block Limit|Eligible|Issued|No Eligible|Active Warps|Registers|Shared Memory Config|Dynamic|Static" | head -30
