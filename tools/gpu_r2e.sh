#!/bin/bash
# Round 2, visit E: persistent Cholesky with the bulk copies issued by all warps (and the slow micro-changes reverted).
mkdir -p gpurun_out
for ctas in 64 96 128; do echo "== $ctas CTAs"; GPIRT_CHOL_CTAS=$ctas timeout 300 python tools/chol_check.py 2>&1 | grep -E "chol_K|backward [0-9.e-]+$|Error|error" | tail -4; done
echo "== per-panel"; GPIRT_CHOL_PERSISTENT=0 timeout 300 python tools/chol_check.py 2>&1 | grep chol_K
echo "== determinism"; timeout 300 python tools/chol_determinism.py 2>&1 | tail -1
echo "== trace n=2000 128"; timeout 120 python tools/chol_trace.py 2000 128 2>&1 | tail -15
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -6
for c in 48 64 96; do
  GPIRT_CHOL_CTAS=$c timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_e_c$c.json 2> gpurun_out/bench_e_c$c.err; echo "bench ctas=$c rc=$?"
done
GPIRT_CHOL_PERSISTENT=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_e_p0.json 2> gpurun_out/bench_e_p0.err
python - <<'PY'
import json
for f in ("bench_e_c48", "bench_e_c64", "bench_e_c96", "bench_e_p0"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 2), "ll", d["ll_after"])
    print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
PY
