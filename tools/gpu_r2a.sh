#!/bin/bash
# Round 2, visit A: parity tests (new cases), smoke, bench with the ESS early rejection on/off, the CPU reference arm once.
set -o pipefail
mkdir -p gpurun_out
nproc; nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -25
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
GPIRT_ESS_EARLY=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_early0.json 2> gpurun_out/bench_early0.err; echo "bench early0 rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_early1.json 2> gpurun_out/bench_early1.err; echo "bench early1 rc=$?"; tail -3 gpurun_out/bench_early1.err
python - <<'PY'
import json
for f in ("gpurun_out/bench_early0.json", "gpurun_out/bench_early1.json"):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 2), "ll", d["ll_after"], "roof", d["roofline"]["frac"], d["roofline"]["peak"])
    print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
    print("  evals", d["mean_evals"], d["proposals_decided_in_fp32"])
    if "cpu_baseline" in d:
        print("  cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["sample"][:160]); print("  tidy", d["cpu_baseline"]["tidy"]["value"])
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; head -c 1500 gpurun_out/bench_ref.json
