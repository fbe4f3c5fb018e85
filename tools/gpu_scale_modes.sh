#!/bin/bash
# bench.py at N = $1 GPUs, once per f* exchange mode (p2p push from the producing kernel / NCCL all-gather)
N=${1:-8}
mkdir -p gpurun_out
for mode in p2p nccl; do
  GPIRT_FSTAR_EXCHANGE=$mode python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline \
    > gpurun_out/bench_n${N}_$mode.json 2> gpurun_out/bench_n${N}_$mode.err
  echo "$mode rc=$?"; tail -2 gpurun_out/bench_n${N}_$mode.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n${N}_$mode.json').read().strip().splitlines()[-1])
print('$mode', {k:d[k] for k in ['value','ms_per_step','n_gpus']}, d['e2e']['value'] if d.get('e2e') else None)
print('   ', ' '.join(f"{k['group']}={k['ms_per_sweep']}" for k in d['kernels'][:12]))
PY
done
