#!/bin/bash
# Round 2, visit C: where the persistent Cholesky spends its time (clock64 trace), reproducibility on small sizes, the flaky constant-IRF test.
mkdir -p gpurun_out
for p in 1 0; do echo "== determinism, persistent=$p"; GPIRT_CHOL_PERSISTENT=$p timeout 300 python tools/chol_determinism.py 2>&1 | tail -7; done
for p in 1 0; do for r in 1 2 3; do echo "== constant-IRF test, persistent=$p run $r"; GPIRT_CHOL_PERSISTENT=$p timeout 300 python -m pytest tests/test_gpu_panel.py -q -k "constant_IRF_runs" 2>&1 | tail -2; done; done
echo "== trace n=2000 128 CTAs"; timeout 120 python tools/chol_trace.py 2000 128 2>&1 | tail -20
echo "== trace n=1001 64 CTAs"; timeout 120 python tools/chol_trace.py 1001 64 2>&1 | tail -20
