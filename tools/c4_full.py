"""BASELINE config 4 in full: synthetic ordinal n = 10 000 respondents x m = 50 000 items, C = 7, one chain item-sharded over the ranks of a
torchrun launch (2 / 4 / 8 B200):
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 tools/c4_full.py [sweeps]
Prints ms per sweep (CUDA events, max over ranks), sweeps/s, the per-group device times of rank 0, and checks that the gathered theta is the
same on every rank and the total log-likelihood finite. The synthetic y is generated per rank from the same seed (every rank keeps its
items' columns and its respondents' rows only)."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpirt_b200.distributed import ShardedSampler  # noqa: E402
from tests.problems import make_problem  # noqa: E402

n, m, C = 10000, 50000, 7
sweeps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
t0 = time.perf_counter()
P = make_problem(n, m, C, seed=20261104 + 4, missing=0.05, dtype="i8")
t_gen = time.perf_counter() - t0
S = ShardedSampler(n, m, C, rank=rank, world=world, device=local, seed=20261104, chain=0)
S.set_data(P["y"])
S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
S.init_draw(draw_beta=False, draw_f=True)
S.sweep(2)
S.sync()
dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
S.sweep(sweeps)
e1.record()
dist.barrier()
torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / sweeps], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
ms = float(t.item())
ll = S.total_ll()
theta = torch.from_numpy(S.get_state(f=False)["theta"]).cuda()
ref = theta.clone()
dist.broadcast(ref, 0)
same = torch.tensor([int(torch.equal(theta, ref))], device="cuda")
dist.all_reduce(same, op=dist.ReduceOp.MIN)
S.profile(True)
S.sweep(1)  # not recorded: it still finds the overlapped pass's shared f* terms valid and would skip those groups
S.profile_reset()
S.sweep(2)
sec = S.profile_read()
S.profile(False)
free, total = torch.cuda.mem_get_info()
if rank == 0:
    ev = {b: round(float(S.diag(b)[0][S.diag(b)[0] > 0].mean()), 2) for b in ("f", "theta", "beta", "thr")}
    print(f"C4 full n={n} m={m} C={C} on {world} GPUs ({m // world} items per GPU): {ms:.2f} ms/sweep = {1e3 / ms:.2f} sweeps/s, ll {ll:.3f}, "
          f"theta identical on all ranks: {bool(same.item())}, mean evals {ev}, device memory in use on rank 0 {(total - free) / 2**30:.1f} GiB, "
          f"y generation {t_gen:.0f} s")
    print("  rank 0 groups (ms/sweep): " + " ".join(f"{k}={v[0] / 2:.2f}" for k, v in sorted(sec.items(), key=lambda kv: -kv[1][0])[:16]))
assert np.isfinite(ll) and bool(same.item())
S.close()
dist.barrier()
dist.destroy_process_group()
