#!/bin/bash
# Round 2, visit Y (2 GPUs): bench lines at N = 1 and N = 2 (both arms at N = 1), smoke.
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/y_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/y_smoke.log
timeout 600 python bench.py > gpurun_out/y_bench1.log 2>&1; echo "bench1 rc=$?"; grep '"metric"' gpurun_out/y_bench1.log | tail -1 | cut -c1-200
timeout 600 python bench.py --impl reference > gpurun_out/y_ref1.log 2>&1; echo "ref rc=$?"; grep '"impl"' gpurun_out/y_ref1.log | tail -1 | cut -c1-700
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 > gpurun_out/y_bench2.log 2>&1; echo "bench2 rc=$?"; grep '"metric"' gpurun_out/y_bench2.log | tail -1 | cut -c1-200
