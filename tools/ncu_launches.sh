#!/bin/bash
# ncu launch list of one warm sweep (after 2 plain sweeps); usage: tools/ncu_launches.sh <skip> <count>
mkdir -p gpurun_out
timeout 300 python tools/sweep_only.py 2 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/sweep_only.py 2 1 > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu.log; wc -l gpurun_out/launches.csv
