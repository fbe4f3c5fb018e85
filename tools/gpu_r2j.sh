#!/bin/bash
# Round 2, visit J: CTA budget of the persistent Cholesky around the default, after L Z / Z* L_post^T became single launches; then the full bench line.
mkdir -p gpurun_out
for c in 32 40 56; do
  GPIRT_CHOL_CTAS=$c timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_j_c$c.json 2> gpurun_out/bench_j_c$c.err; echo "bench ctas=$c rc=$?"
done
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_j_default.json 2> gpurun_out/bench_j_default.err; echo "bench default rc=$?"
python - <<'PY'
import json
for f in ("bench_j_c32", "bench_j_c40", "bench_j_c56", "bench_j_default"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 2), "launches", d["gpu_launches"], "roof", d["roofline"]["frac"], "loglik", round(d["loglik_gbs"], 1))
    print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
    if "cpu_baseline" in d: print("  cpu", d["cpu_baseline"]["value"], "tidy", d["cpu_baseline"]["tidy"]["value"], "chains", d["batched_chains"]["value"])
PY
