#!/bin/bash
# Round 2, visit O (2 GPUs): the distributed theta-only chain — parity tests, the one-process check at n = 10000, config 4 halves on 2 GPUs.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/o_multi.log 2>&1; echo "multi rc=$?"; tail -5 gpurun_out/o_multi.log
timeout 600 python tools/dist_chain_check.py 10000 64 3 2 > gpurun_out/o_check.log 2>&1; echo "check rc=$?"; tail -6 gpurun_out/o_check.log
GPIRT_DIST_CHAIN_SAME_DEVICE=1 CUDA_VISIBLE_DEVICES=0 timeout 300 python - > gpurun_out/o_same.log 2>&1 <<'PY'
import os, sys
sys.path.insert(0, ".")
import numpy as np, gpirt_b200
from gpirt_b200.distributed import LocalShards
from tests.problems import make_problem
n, m, C = 4203, 21, 4
P = make_problem(n, m, C, seed=5)
def run(S):
    S.set_data(P["y"]); S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"]); S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True); S.sweep(3); S.sync(); st = S.get_state(fstar=False); st["ll"] = S.total_ll(); S.close(); return st
a = run(gpirt_b200.Sampler(n, m, C, seed=77))
for w in (2, 3):
    L = LocalShards(n, m, C, devices=[0] * w, seed=77); print("dist", L.dist_chain)
    b = run(L)
    print("SAME_DEVICE", w, all(np.array_equal(a[k], b[k]) for k in ("theta", "beta", "thresholds", "f")) and a["ll"] == b["ll"], flush=True)
PY
echo "same-device rc=$?"; tail -5 gpurun_out/o_same.log
for d in 1 0; do
GPIRT_DIST_CHAIN=$d timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/c4_full.py 4 > gpurun_out/o_c4_dist$d.log 2>&1
echo "c4 dist=$d rc=$?"; grep -E "C4 full|rank 0 groups" gpurun_out/o_c4_dist$d.log || tail -8 gpurun_out/o_c4_dist$d.log
done
