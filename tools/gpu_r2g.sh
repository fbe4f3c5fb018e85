#!/bin/bash
# Round 2, visit G (2 GPUs): single-GPU sharding tests, 2-GPU identity tests, bench at N = 1 and 2 in both theta-step schemes.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shards_local.py -q -x 2>&1 | tail -15
timeout 900 python -m pytest tests/test_gpu_multi.py -q 2>&1 | tail -15
for mode in items respondents; do
  GPIRT_THETA_SHARD=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_g_2gpu_$mode.json 2> gpurun_out/bench_g_2gpu_$mode.err; echo "bench 2gpu $mode rc=$?"; tail -3 gpurun_out/bench_g_2gpu_$mode.err
done
python - <<'PY'
import json
for f in ("bench_g_2gpu_items", "bench_g_2gpu_respondents"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 2), "ll", d["ll_after"], "same", d.get("sharded_equals_single"))
    print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
    print("  ", d["config"]["parallelism"], d["mean_evals"])
PY
