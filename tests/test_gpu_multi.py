"""Item-sharded sweeps on 2 GPUs (one process per GPU, NCCL) against the same chain on 1 GPU: bit-identical state, because the
Philox streams are keyed by global item / respondent index and every reduction has a fixed order. Skipped with fewer than 2 GPUs."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, os.environ["GPIRT_ROOT"])
import gpirt_b200
from gpirt_b200.distributed import ShardedSampler
from tests.problems import make_problem
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
n, m, C, sweeps = 203, 61, 4, 3   # ragged on purpose: 61 items -> chunks of 31 + 30, 203 respondents -> 102 + 101
if os.environ.get("GPIRT_TEST_BIG"):  # config-4 respondent counts: the theta-only chain distributed over the ranks (two-level Cholesky, n >= 4096)
    n, m, sweeps = 4203, 21, 2
P = make_problem(n, m, C, seed=5)
def run(S):
    S.set_data(P["y"]); S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True)
    S.sweep(sweeps); S.sync()
    st = S.get_state(fstar=False); st["ll"] = S.total_ll()
    S.close()
    return st
S = ShardedSampler(n, m, C, rank=rank, world=world, device=local, seed=77)
if rank == 0:
    print(f"DIST_CHAIN {S.dist_chain}", flush=True)
sh = run(S)
# gather the f columns of every rank on rank 0 (get_state writes only the local item block)
f = torch.from_numpy(np.ascontiguousarray(sh["f"].T)).cuda()
dist.all_reduce(f)  # the other ranks' blocks are zero
if rank == 0:
    one = run(gpirt_b200.Sampler(n, m, C, seed=77, device=local))
    ok = all(np.array_equal(sh[k], one[k]) for k in ("theta", "beta", "thresholds")) and sh["ll"] == one["ll"]
    ok = ok and np.array_equal(f.cpu().numpy().T, one["f"])
    print("WORLD2_EQUALS_WORLD1", ok, sh["ll"], one["ll"])
# the sharded mirror of gpirtMCMC (partial y uploads, device-side draw store) against the single-GPU entry point
from gpirt_b200.distributed import gpirtMCMC_sharded
kw = dict(beta_prior_sds=P["beta_prior_sds"], theta_prior_sds=P["theta_prior_sds"], theta_init=P["theta0"], thresholds=P["thr0"], SEED=123, beta_init=P["beta0"])
a = gpirtMCMC_sharded(P["y"], 5 if n < 1000 else 3, 1, THIN=2, rank=rank, world=world, device=local, **kw)
if rank == 0:
    b = gpirt_b200.gpirtMCMC(P["y"], 5 if n < 1000 else 3, 1, THIN=2, CHAIN=1, device=local, **kw)[0]
    ok = np.array_equal(a["theta"], b["theta"]) and np.array_equal(a["ll"], b["ll"]) and len(a["beta"]) == len(b["beta"]) == (2 if n < 1000 else 1)
    ok = ok and all(np.array_equal(x, y) for x, y in zip(a["beta"] + a["threshold"], b["beta"] + b["threshold"]))
    print("SHARDED_MCMC_EQUALS_SINGLE", ok, a["ll"], b["ll"])
dist.barrier(); dist.destroy_process_group()
'''


def _gpus():
    from gpirt_b200 import _lib
    return _lib.lib().gpirt_device_count()


# items: the theta step sharded by items, partial sums exchanged round by round (default); respondents-p2p / respondents-nccl: the theta step sharded
# by respondents, f* chunks pushed through NVLink peer memory by the producing kernel / NCCL all-gather
# items-dist: config-4 respondent counts, where the theta-only chain (Cholesky of K, L^-1 k*, K^-1 k*) is computed once across the ranks
@pytest.mark.parametrize("exchange", ["items", "p2p", "nccl", "items-dist"])
def test_two_gpu_chain_is_bit_identical_to_one_gpu(tmp_path, exchange):
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    items = exchange.startswith("items")
    env = dict(os.environ, GPIRT_ROOT=ROOT, GPIRT_FSTAR_EXCHANGE=exchange if not items else "nccl", GPIRT_THETA_SHARD="items" if items else "respondents")
    if exchange == "items-dist":
        env["GPIRT_TEST_BIG"] = "1"
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), str(script)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "WORLD2_EQUALS_WORLD1 True" in r.stdout, r.stdout[-2000:]
    assert "SHARDED_MCMC_EQUALS_SINGLE True" in r.stdout, r.stdout[-2000:]
    assert ("DIST_CHAIN True" in r.stdout) == (exchange == "items-dist")


@pytest.mark.parametrize("n,m,iters", [(203, 61, 5), (4203, 21, 3)])  # the second: the theta-only chain distributed over the two devices
def test_gpirtMCMC_over_two_gpus_from_one_process_equals_one_gpu(n, m, iters):
    """gpirt_mcmc(n_devices = 2) with two real devices: peer copies over NVLink, no process group."""
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import gpirt_b200
    from tests.problems import make_problem

    C = 4
    P = make_problem(n, m, C, seed=5)
    kw = dict(beta_prior_sds=P["beta_prior_sds"], theta_prior_sds=P["theta_prior_sds"], theta_init=P["theta0"], thresholds=P["thr0"], SEED=123, beta_init=P["beta0"])
    a = gpirt_b200.gpirtMCMC(P["y"], iters, 1, THIN=2, devices=[0, 1], **kw)[0]
    b = gpirt_b200.gpirtMCMC(P["y"], iters, 1, THIN=2, **kw)[0]
    assert np.array_equal(a["theta"], b["theta"]) and np.array_equal(a["ll"], b["ll"])
    assert all(np.array_equal(x, y) for x, y in zip(a["beta"] + a["threshold"], b["beta"] + b["threshold"]))
