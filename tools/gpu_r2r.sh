#!/bin/bash
# Round 2, visit R (1 GPU): band tile order of the fp64 GEMM — full GPU suite, bench with the order on / off, ncu --set full of the L*Z launch.
mkdir -p gpurun_out /tmp/ncu
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r_tests.log
timeout 600 python bench.py > gpurun_out/r_bench.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/r_bench.log | tail -1 | cut -c1-400
GPIRT_GEMM_BAND_MB=0 timeout 600 python bench.py > gpurun_out/r_bench_band0.log 2>&1; echo "bench band0 rc=$?"; grep '"metric"' gpurun_out/r_bench_band0.log | tail -1 | cut -c1-400
timeout 300 python tools/sweep_only.py 5 2 > gpurun_out/r_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_f64_kernel" --launch-count 8 -f -o /tmp/ncu/r02_gemm_band python tools/sweep_only.py 5 2 > gpurun_out/r_ncu.log 2>&1
echo "ncu rc=$?"; tail -1 gpurun_out/r_ncu.log
python tools/ncu_summary.py /tmp/ncu/r02_gemm_band.ncu-rep > gpurun_out/r02_ncu_gemm_band.txt 2> gpurun_out/r_ncu_summary.err; grep -n "==\|duration\|grid  \|dram\|DMMA" gpurun_out/r02_ncu_gemm_band.txt | head -60
