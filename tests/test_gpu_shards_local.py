"""The item-sharded chain driven from one process on ONE GPU (gpirt_b200.distributed.LocalShards: several shards of one chain on the same device,
exchanges as device copies): the sharded logic — partition, the item-sharded theta rounds, the packed exchange — against the CPU oracle and
against the single-handle chain, on the box the driver runs (tests/test_gpu_multi.py needs two GPUs and NCCL and is skipped there).
Reference: the per-respondent loop of src/draw-theta.cpp:121-170 and the per-item loops of src/draw-f.cpp:68, src/draw-beta.cpp:64."""
import numpy as np
import pytest

import gpirt_b200
from gpirt_b200.distributed import LocalShards
from oracle import orc
from tests.problems import make_problem, uniform_streams

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,m,C,world", [(150, 91, 5, 3), (64, 1500, 2, 2), (90, 40, 4, 4)])
def test_item_sharded_theta_step_takes_the_oracles_decisions(n, m, C, world):
    """draw_theta with supplied normals / uniforms: every rank scores all respondents against its own items, partial sums exchanged round by
    round. theta bit-exact, identical proposal counts (src/draw-theta.cpp:26-84); m = 1500 needs more than one round of 16 proposals."""
    P = make_problem(n, m, C, seed=31, on_grid=True)
    rng = P["rng"]
    G = LocalShards(n, m, C, devices=[0] * world, seed=1)
    G.set_data(P["y"])
    G.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    O = orc.Sampler(P["y"], C, P["beta_prior_sds"], P["theta_prior_sds"])
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    fstar = np.asfortranarray(rng.standard_normal((orc.NGRID, m)) * (0.3 if m > 1000 else 1.0))
    beta = P["beta0"] * (0.1 if m > 1000 else 1.0)
    G.set_state(theta=P["theta0"], beta=beta, f=f0, thresholds=P["thr0"], fstar=fstar)
    O.set_state(theta=P["theta0"], beta=beta, f=f0, thr=P["thr0"], fstar=fstar)
    O.compute_mu()
    z = np.asfortranarray(rng.standard_normal((1, n)))
    u = uniform_streams(rng, n, stride=256)
    if m > 1000:
        u[1:, :] = 0.9  # brackets that shrink by a tenth per rejection: up to 40 proposals per respondent, i.e. three exchange rounds
    G.draw_theta(z=z, u=u)
    O.draw_theta(orc.supplied(z=z, u=u))
    ev_o, mg_o, over = O.diag("theta")
    assert not over and mg_o.min() > 1e-9
    for S in G.S:  # every shard holds the whole result
        assert np.array_equal(S.diag("theta")[0], ev_o)
        th = np.empty(n)
        ptr, nbytes = S.exchange_buffer("theta")
        import torch

        from gpirt_b200.distributed import _DevBuf

        th[:] = torch.as_tensor(_DevBuf(ptr, nbytes * world), device="cuda:0")[:n].cpu().numpy()
        assert np.array_equal(th, O.state()["theta"])
    assert G.theta_rounds == (1 if ev_o.max() - 1 <= 16 else -(-(int(ev_o.max()) - 1) // 16))
    if m > 1000:
        assert G.theta_rounds >= 2
    G.close()


@pytest.mark.parametrize("theta_shard", ["items", "respondents"])
@pytest.mark.parametrize("n,m,C,world", [(320, 131, 5, 2), (200, 67, 3, 3)])
def test_sharded_chain_on_one_gpu_equals_the_single_handle_chain(n, m, C, world, theta_shard):
    """Production sweeps (Philox streams keyed by global unit): ragged item shards, both theta-step schemes, bit for bit against one handle."""
    P = make_problem(n, m, C, seed=99, missing=0.07)
    out = []
    for sharded in (False, True):
        kw = dict(seed=7, chain=0)
        S = LocalShards(n, m, C, devices=[0] * world, theta_shard=theta_shard, **kw) if sharded else gpirt_b200.Sampler(n, m, C, **kw)
        S.set_data(P["y"])
        S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
        S.init_draw(draw_beta=False, draw_f=True)
        S.sweep(3)
        S.sync()
        st = S.get_state()
        st["ll"] = S.total_ll()
        st["ev"] = {b: S.diag(b)[0] for b in ("f", "theta", "beta", "thr")}
        out.append(st)
        S.close()
    a, b = out
    for k in ("theta", "beta", "thresholds", "f"):
        assert np.array_equal(a[k], b[k]), k
    for blk in ("f", "theta", "beta", "thr"):
        assert np.array_equal(a["ev"][blk], b["ev"][blk]), blk
    assert a["ll"] == b["ll"]
