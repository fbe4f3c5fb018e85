#!/bin/bash
# bench.py at N = $1 GPUs under torchrun (as the driver launches it), JSON line -> gpurun_out/bench_n$1.json
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline \
  > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$? lines=$(wc -l < gpurun_out/bench_n$N.json)"; tail -2 gpurun_out/bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ['value','ms_per_step','n_gpus']}, d['e2e']['value'] if d.get('e2e') else None)
PY
