#!/bin/bash
# bench.py at N GPUs (C3 weak scaling, item-sharded theta step): tools/gpu_scale_n.sh N
N=${1:-4}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_s_${N}gpu.json 2> gpurun_out/bench_s_${N}gpu.err
echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|NCCL version" gpurun_out/bench_s_${N}gpu.err | tail -3
python - <<PY
import json
d = json.load(open("gpurun_out/bench_s_${N}gpu.json"))
print("N=$N value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "same", d.get("sharded_equals_single"), "evals", d["mean_evals"]["theta"])
print("  ", {k["group"]: k["ms_per_sweep"] for k in d["kernels"]})
PY
