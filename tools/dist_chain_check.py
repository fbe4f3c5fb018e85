"""The distributed theta-only chain from ONE process (LocalShards over devices 0..W-1, chain buffers attached without IPC) against the same chain on one
GPU at a config-4 respondent count: same bits, and the time per sweep with the chain distributed / replicated.
   python tools/dist_chain_check.py [n] [m] [sweeps] [world]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200  # noqa: E402
from gpirt_b200.distributed import LocalShards  # noqa: E402
from tests.problems import make_problem  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 64
sweeps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
world = int(sys.argv[4]) if len(sys.argv) > 4 else 2
C = 5
P = make_problem(n, m, C, seed=9)


def run(S, label):
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True)
    S.sweep(1)
    S.sync()
    t0 = time.perf_counter()
    S.sweep(sweeps)
    S.sync()
    ms = (time.perf_counter() - t0) * 1e3 / sweeps
    st = S.get_state(fstar=False)
    st["ll"] = S.total_ll()
    print(f"{label}: {ms:.2f} ms/sweep, ll {st['ll']:.6f}", flush=True)
    # the overlapped schedule's own timeline (profiling level 2: section times include waiting for peers and sharing the SMs)
    each = (lambda f, *a: [getattr(x, f)(*a) for x in S.S]) if hasattr(S, "S") and isinstance(S.S, list) else (lambda f, *a: [getattr(S, f)(*a)])
    each("profile", 2)
    each("profile_reset")
    S.sweep(2)
    S.sync()
    sec = each("profile_read")[0]
    each("profile", 0)
    print("   shard 0 sections (ms/sweep): " + " ".join(f"{k}={v[0] / 2:.2f}" for k, v in sorted(sec.items(), key=lambda kv: -kv[1][0])[:12]), flush=True)
    S.close()
    return st


one = run(gpirt_b200.Sampler(n, m, C, seed=77, device=0), "one GPU")
os.environ["GPIRT_DIST_CHAIN"] = "0"
rep = run(LocalShards(n, m, C, devices=list(range(world)), seed=77), f"{world} shards, chain replicated")
os.environ["GPIRT_DIST_CHAIN"] = "1"
L = LocalShards(n, m, C, devices=list(range(world)), seed=77)
assert L.dist_chain
dis = run(L, f"{world} shards, chain distributed")
for name, st in (("replicated", rep), ("distributed", dis)):
    ok = all(np.array_equal(st[k], one[k]) for k in ("theta", "beta", "thresholds", "f")) and st["ll"] == one["ll"]
    print(f"DIST_CHAIN_CHECK {name} equals one GPU: {ok}")
