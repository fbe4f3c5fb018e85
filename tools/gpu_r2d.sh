#!/bin/bash
# Round 2, visit D: persistent Cholesky after the trace-driven changes (priority CTAs, L D L^T pivot chain, batched mma phases, symmetric updates).
mkdir -p gpurun_out
echo "== check + timing, default CTAs"; timeout 300 python tools/chol_check.py 2>&1 | tail -19
for ctas in 48 96 132; do echo "== $ctas CTAs"; GPIRT_CHOL_CTAS=$ctas timeout 300 python tools/chol_check.py 2>&1 | grep chol_K; done
echo "== determinism"; timeout 300 python tools/chol_determinism.py 2>&1 | tail -6
echo "== trace n=2000 128"; timeout 120 python tools/chol_trace.py 2000 128 2>&1 | tail -16
echo "== trace n=1001 64"; timeout 120 python tools/chol_trace.py 1001 64 2>&1 | tail -16
echo "== constant-IRF debug persistent=1"; timeout 120 python tools/cst_debug.py 2.0 2>&1 | tail -20
echo "== constant-IRF debug persistent=0"; GPIRT_CHOL_PERSISTENT=0 timeout 120 python tools/cst_debug.py 2.0 2>&1 | tail -10
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -8
for c in 64 96 128; do
  GPIRT_CHOL_CTAS=$c timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_d_c$c.json 2> gpurun_out/bench_d_c$c.err; echo "bench ctas=$c rc=$?"
done
python - <<'PY'
import json
for f in ("bench_d_c64", "bench_d_c96", "bench_d_c128"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 2), "ll", d["ll_after"])
    print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
PY
