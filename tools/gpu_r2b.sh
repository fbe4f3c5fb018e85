#!/bin/bash
# Round 2, visit B: the persistent Cholesky — sanitizer on small sizes, correctness + timing against the per-panel path, full tests, bench A/B.
set -o pipefail
mkdir -p gpurun_out
echo "== racecheck (small sizes)"; timeout 600 compute-sanitizer --tool racecheck --print-limit 5 python tools/chol_check.py quick 2>&1 | tail -15
echo "== memcheck (small sizes)"; timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python tools/chol_check.py quick 2>&1 | tail -8
for ctas in 32 64 96 128; do echo "== persistent, $ctas CTAs"; GPIRT_CHOL_CTAS=$ctas timeout 300 python tools/chol_check.py 2>&1 | tail -6; done
echo "== per-panel path"; GPIRT_CHOL_PERSISTENT=0 timeout 300 python tools/chol_check.py 2>&1 | tail -4
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -15
for p in 1 0; do
  GPIRT_CHOL_PERSISTENT=$p timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_chol$p.json 2> gpurun_out/bench_chol$p.err; echo "bench persistent=$p rc=$?"
done
GPIRT_CHOL_CTAS=32 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_chol1_c32.json 2> gpurun_out/bench_chol1_c32.err
GPIRT_CHOL_CTAS=112 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_chol1_c112.json 2> gpurun_out/bench_chol1_c112.err
python - <<'PY'
import json
for f in ("bench_chol1", "bench_chol0", "bench_chol1_c32", "bench_chol1_c112"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 2), "ll", d["ll_after"])
    print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
PY
