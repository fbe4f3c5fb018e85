#!/bin/bash
mkdir -p gpurun_out
for b in 0 12 24 41 82; do GPIRT_GEMM_BAND_MB=$b timeout 200 python tools/band_sweep.py 2>&1 | tail -1; done | tee gpurun_out/t_band.log
