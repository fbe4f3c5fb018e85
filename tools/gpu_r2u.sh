#!/bin/bash
# Round 2, visit U (1 GPU): second half of the theta-only chain deferred behind L*Z (on / off), band order at 41 MB (ncu traffic of L*Z), tests.
mkdir -p gpurun_out /tmp/ncu
for d in 1 0 1 0; do GPIRT_DEFER_PREP=$d timeout 200 python tools/band_sweep.py 2>&1 | tail -1 | sed "s/^/defer=$d /"; done | tee gpurun_out/u_defer.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/u_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/u_tests.log
timeout 600 python bench.py > gpurun_out/u_bench.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/u_bench.log | tail -1 | cut -c1-260
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:gemm_f64_kernel" --launch-count 2 -f -o /tmp/ncu/r02_gemm_band41 python tools/sweep_only.py 1 1 > gpurun_out/u_ncu.log 2>&1
echo "ncu rc=$?"; python tools/ncu_summary.py /tmp/ncu/r02_gemm_band41.ncu-rep > gpurun_out/r02_ncu_gemm_band41.txt 2> gpurun_out/u_ncu_summary.err; grep -n "==\|duration\|grid  \|dram\|DMMA" gpurun_out/r02_ncu_gemm_band41.txt | head -12
