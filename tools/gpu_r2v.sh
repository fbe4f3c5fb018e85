#!/bin/bash
# Round 2, visit V (1 GPU): CTAs of the persistent Cholesky inside a sweep, re-tuned after the ESS kernels got faster.
mkdir -p gpurun_out
for c in 32 48 64 80 96 128; do GPIRT_CHOL_CTAS=$c timeout 200 python tools/band_sweep.py 2>&1 | tail -1 | sed "s/^/chol_ctas=$c /"; done | tee gpurun_out/v_ctas.log
for t in 512 768 1024; do GPIRT_ESS_THREADS_PER_SM=$t timeout 200 python tools/band_sweep.py 2>&1 | tail -1 | sed "s/^/ess_threads_per_sm=$t /"; done | tee -a gpurun_out/v_ctas.log
