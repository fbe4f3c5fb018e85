#!/bin/bash
# Round 2, visit P (2 GPUs): distributed chain with the split trtri levels — parity, section timeline at n = 10000, config 4 on 2 GPUs.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/p_multi.log 2>&1; echo "multi rc=$?"; tail -5 gpurun_out/p_multi.log
timeout 600 python tools/dist_chain_check.py 10000 64 3 2 > gpurun_out/p_check.log 2>&1; echo "check rc=$?"; tail -9 gpurun_out/p_check.log
GPIRT_DIST_CHAIN=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/c4_full.py 4 > gpurun_out/p_c4_dist1.log 2>&1
echo "c4 dist=1 rc=$?"; grep -E "C4 full|rank 0 groups" gpurun_out/p_c4_dist1.log || tail -8 gpurun_out/p_c4_dist1.log
