#!/bin/bash
# Round 2, visit H (8 GPUs): C3 weak scaling with the item-sharded theta step against the respondent-sharded one, and config 4 through bench.py.
mkdir -p gpurun_out
run() {  # name, env, extra args
  env $2 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 $3 > gpurun_out/bench_h_$1.json 2> gpurun_out/bench_h_$1.err
  echo "bench $1 rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|NCCL version" gpurun_out/bench_h_$1.err | tail -4
}
run c3_items GPIRT_THETA_SHARD=items "--steps 20 --warmup 5"
run c3_respondents GPIRT_THETA_SHARD=respondents "--steps 20 --warmup 5"
run c4_items GPIRT_THETA_SHARD=items "--config c4 --steps 10 --warmup 3"
python - <<'PY'
import json
for f in ("c3_items", "c3_respondents", "c4_items"):
    try:
        d = json.load(open(f"gpurun_out/bench_h_{f}.json"))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 2), "ll", d["ll_after"], "same", d.get("sharded_equals_single"))
    print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
    print("  ", d["mean_evals"], d["config"]["parallelism"][-120:])
PY
