#!/bin/bash
# Round 2, visit Q (8 GPUs): config 4 with the theta-only chain distributed / replicated, same box; then the bench line.
mkdir -p gpurun_out
for d in 1 0; do
GPIRT_DIST_CHAIN=$d timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/c4_full.py 6 > gpurun_out/q_c4_dist$d.log 2>&1
echo "c4 dist=$d rc=$?"; grep -E "C4 full|rank 0 groups" gpurun_out/q_c4_dist$d.log || tail -8 gpurun_out/q_c4_dist$d.log
done
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --config c4 --steps 6 --warmup 3 > gpurun_out/q_bench_c4.log 2>&1
echo "bench rc=$?"; grep '"metric"' gpurun_out/q_bench_c4.log | tail -1 | cut -c1-1500
