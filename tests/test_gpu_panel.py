"""horizon > 1 (dynamic GPIRT): the CUDA panel (gpirt_panel_*: H single-session samplers + the joint theta step) against the CPU oracle's
panel, which is pinned bit for bit to the reference's own gpirtMCMC at H = 3 (tests/test_oracle_vs_ref.py). theta is bit-exact (it lives on
the 0.01 grid), proposal counts identical, beta / thresholds 1e-9, f / f* norm-wise 1e-6 on well-conditioned priors."""
import numpy as np
import pytest

from oracle import orc
from tests.problems import make_problem, uniform_streams

pytestmark = pytest.mark.gpu
REGIMES = [("RDM", 0.05, "Matern"), ("GP", 2.0, "Matern"), ("GP", 4.5, "RBF")]


def problem(n, m, C, H, seed):
    Ps = [make_problem(n, m, C, seed=seed + h, missing=0.12, on_grid=True) for h in range(H)]
    rng = np.random.default_rng(seed + 100)
    P = dict(n=n, m=m, C=C, H=H, rng=rng)
    P["y"] = np.asfortranarray(np.stack([p["y"] for p in Ps], axis=2))
    P["theta0"] = np.asfortranarray(np.stack([p["theta0"] for p in Ps], axis=1))
    P["beta0"] = np.asfortranarray(np.stack([p["beta0"] * 0.3 for p in Ps], axis=2))
    thr = []
    for h in range(H):  # different thresholds per session: the RDM step scores every session against session 0's
        t = Ps[h]["thr0"].copy()
        t[:, 1:-1] += 0.15 * h + 0.05 * rng.standard_normal((m, 1))
        thr.append(t)
    P["thr0"] = np.asfortranarray(np.stack(thr, axis=2))
    P["bps"] = np.asfortranarray(np.tile(np.array([[0.8], [0.5], [0.1]]), (1, m)))  # well-conditioned K
    P["tps"] = np.asfortranarray(np.vstack([rng.uniform(0.0, 0.8, (1, n)), np.zeros((1, n))]))
    return P


def make_pair(P, ls, kernel, os_=1.3, seed=99):
    import gpirt_b200 as g

    G = g.Panel(P["n"], P["m"], P["C"], P["H"], theta_os=os_, theta_ls=ls, KERNEL=kernel, seed=seed, chain=2)
    G.set_data(P["y"])
    G.set_priors(P["bps"], P["tps"])
    O = orc.Panel(P["y"], P["C"], P["bps"], P["tps"], theta_os=os_, theta_ls=ls, KERNEL=kernel)
    return G, O


@pytest.mark.parametrize("regime,ls,kernel", REGIMES)
def test_panel_theta_step_same_states_given_same_streams(regime, ls, kernel):
    n, m, C, H = 90, 26, 4, 3
    P = problem(n, m, C, H, seed=3)
    rng = P["rng"]
    G, O = make_pair(P, ls, kernel)
    assert G.regime == regime
    f = np.asfortranarray(rng.standard_normal((n, m, H)))
    fstar = np.asfortranarray(rng.standard_normal((orc.NGRID, m, H)) * 0.7)
    G.set_state(theta=P["theta0"], beta=P["beta0"], f=f, thresholds=P["thr0"], fstar=fstar)
    O.set_state(theta=P["theta0"], beta=P["beta0"], f=f, thr=P["thr0"], fstar=fstar)
    O.compute_mu()
    if regime == "GP":
        assert np.max(np.abs(G.L_time() - O.L_time)) < 1e-15
        z, u = np.asfortranarray(rng.standard_normal((H, n))), uniform_streams(rng, n)
        G.draw_theta(z=z, u=u)
        O.draw_theta(orc.supplied(z=z, u=u))
        ev_g, mg = G.diag_theta()
        ev_o = O.S[0].diag("theta")[0]
    else:
        z = [np.asfortranarray(rng.standard_normal((1, n))) for _ in range(H)]
        u = [uniform_streams(rng, n) for _ in range(H)]
        G.draw_theta(z=z, u=u)
        O.draw_theta([orc.supplied(z=z[h], u=u[h]) for h in range(H)])
        ev_g, mg = G.diag_theta()
        ev_o = np.stack([S.diag("theta")[0] for S in O.S])
    assert np.array_equal(ev_g, ev_o), "different numbers of proposals"
    assert np.min(mg) > 1e-9  # no accept decision sat on a knife edge: the comparison below is meaningful
    assert np.array_equal(G.get_state()["theta"], O.state()["theta"])
    assert np.any(ev_g > 2)


@pytest.mark.parametrize("regime,ls,kernel", REGIMES)
def test_panel_sweeps_match_oracle_on_the_same_philox_streams(regime, ls, kernel):
    n, m, C, H, sweeps = 70, 18, 4, 3, 3
    P = problem(n, m, C, H, seed=11)
    G, O = make_pair(P, ls, kernel, seed=4242)
    rng = P["rng"]
    f0 = np.asfortranarray(rng.standard_normal((n, m, H)) * 0.5)
    G.set_state(theta=P["theta0"], beta=P["beta0"], f=f0, thresholds=P["thr0"])
    O.set_state(theta=P["theta0"], beta=P["beta0"], f=f0, thr=P["thr0"])
    for it in range(sweeps):
        G.set_iter(it)
        G.sweep(1)
        assert O.sweep(lambda block, h: orc.philox(4242, 2, it, session=0 if h is None else h)) == 0
        a, b = G.get_state(fstar=True), O.state()
        assert np.array_equal(a["theta"], b["theta"]), f"sweep {it}: theta"
        assert np.max(np.abs(a["beta"] - b["beta"])) < 1e-9 and np.max(np.abs(a["thresholds"][:, 1:-1] - b["thr"][:, 1:-1])) < 1e-9
        assert np.max(np.abs(a["f"] - b["f"])) < 1e-6 * max(1.0, np.max(np.abs(b["f"])))
        assert np.max(np.abs(a["fstar"] - b["fstar"])) < 1e-6 * max(1.0, np.max(np.abs(b["fstar"])))
        assert abs(G.total_ll() - O.total_ll()) <= 1e-9 * abs(O.total_ll())


def test_panel_refuses_the_regime_the_reference_aborts_in():
    import gpirt_b200 as g

    with pytest.raises(g.GpirtError) as e:
        g.Panel(20, 5, 3, 2, theta_ls=10.0)  # the R default theta_ls = 10 >= 3H for H <= 3
    assert "out of bounds" in str(e.value) or "thresholds" in str(e.value)
    P1 = g.Panel(20, 5, 3, 1, theta_ls=10.0)  # horizon 1 is the plain sampler
    assert P1.regime == "H1"


@pytest.mark.parametrize("regime,ls,kernel", REGIMES[:2])
def test_gpirtMCMC_at_horizon_3_equals_the_stepwise_panel(regime, ls, kernel):
    """gpirt_mcmc with horizon > 1 (cubes in, cubes out, burn-in / thinning, per-session device stores flushed slice by slice) against the same
    chain driven through the panel calls: bit for bit."""
    import gpirt_b200 as g

    n, m, C, H, sample, burn, thin = 60, 14, 4, 3, 6, 1, 2
    P = problem(n, m, C, H, seed=21)
    res = g.gpirtMCMC(P["y"], sample, burn, THIN=thin, CHAIN=1, beta_prior_sds=P["bps"], theta_prior_sds=P["tps"], theta_os=1.3, theta_ls=ls,
                      KERNEL=kernel, theta_init=P["theta0"], thresholds=P["thr0"], SEED=606, beta_init=P["beta0"], store_f=True, store_fstar=True)[0]
    ns = sample // thin
    assert res["theta"].shape == (ns, n, H) and res["beta"][0].shape == (3, m, H) and res["fstar"][0].shape == (g.NGRID, m, H)
    G = g.Panel(n, m, C, H, theta_os=1.3, theta_ls=ls, KERNEL=kernel, seed=606, chain=0)
    G.set_data(P["y"])
    G.set_priors(P["bps"], P["tps"])
    G.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    G.set_iter(0xFFFFFFFF)
    for S in G.S:
        S.init_draw(draw_beta=False, draw_f=True)
    G.set_iter(0)
    s = 0
    for it in range(sample + burn):
        G.sweep(1)
        if it >= burn and it % thin == 0 and (it - burn) // thin < ns:
            st = G.get_state(fstar=True)
            assert np.array_equal(res["theta"][s], st["theta"]) and np.array_equal(res["beta"][s], st["beta"])
            assert np.array_equal(res["threshold"][s], st["thresholds"]) and np.array_equal(res["f"][s], st["f"])
            assert np.array_equal(res["fstar"][s], st["fstar"])
            assert res["ll"][s] == G.total_ll()
            s += 1
    assert s == ns


@pytest.mark.parametrize("regime,ls", [("RDM", 0.05), ("GP", 2.0)])
def test_gpirtMCMC_constant_IRF_runs_the_constant_panel(regime, ls):
    """gpirtMCMC(constant_IRF = 1) through the C entry point (gpirt_mcmc -> mcmc_constant, csrc/entry.cu): one f* and one set of thresholds for all
    sessions in every kept draw, per-session beta, finite log-likelihood; bit-identical to the block-by-block ConstantPanel loop in Python
    (constant_panel.py, the composition the golden replays pin to the reference); another chain id gives a different chain."""
    import gpirt_b200 as g
    from gpirt_b200 import _lib

    n, m, C, H = 60, 9, 4, 3
    Ps = [make_problem(n, m, C, seed=300 + h, missing=0.1) for h in range(H)]
    y = np.asfortranarray(np.stack([P["y"] for P in Ps], axis=2))
    th0 = np.asfortranarray(np.stack([Ps[0]["theta0"]] * H, axis=1))
    kw = dict(theta_ls=ls, theta_os=1.2, theta_init=th0, thresholds=Ps[0]["thr0"], SEED=9, constant_IRF=1, store_f=True, store_fstar=True)
    a = g.gpirtMCMC(y, 4, 2, THIN=2, **kw)[0]
    assert a["theta"].shape == (2, n, H) and len(a["beta"]) == 2 and np.all(np.isfinite(a["ll"])) and np.all(a["ll"] < 0)
    for k in range(2):
        fs, thr, be = a["fstar"][k], a["threshold"][k], a["beta"][k]
        assert fs.shape == (1001, m, H) and all(np.array_equal(fs[:, :, 0], fs[:, :, h]) for h in range(1, H))
        assert all(np.array_equal(thr[:, :, 0], thr[:, :, h]) for h in range(1, H))
        assert not np.array_equal(be[:, :, 0], be[:, :, 1])
        assert np.all(np.abs(a["theta"][k]) <= 5.0)
    from gpirt_b200.api import _as_cube, _gpirtMCMC_constant_irf

    b = _gpirtMCMC_constant_irf(_as_cube(y), 4, 2, 2, 1, None, None, None, None, 1.2, ls, "Matern", th0, Ps[0]["thr0"], 9, True, True, 0, None, 0)[0]
    assert np.array_equal(a["theta"], b["theta"]) and np.array_equal(a["ll"], b["ll"])
    for k in range(2):
        for key in ("beta", "threshold", "f", "fstar"):
            assert np.array_equal(a[key][k], b[key][k]), f"{key} draw {k}: C entry point != stepwise ConstantPanel"
    c = g.gpirtMCMC(y, 4, 2, THIN=2, first_chain=1, **kw)[0]
    assert not np.array_equal(a["theta"], c["theta"])


@pytest.mark.gpu
def test_gpirtMCMC_constant_IRF_with_one_session():
    """constant_IRF = 1 at horizon 1 with R's default theta_ls = 10 (CST regime): a legal call in the reference (the f* draw then runs through the
    100 inducing points); C entry point == the block-by-block ConstantPanel loop, bit for bit. The composition itself is pinned to the reference's
    own run by tests/golden/cst_cst_h1.npz (test_golden.py)."""
    import gpirt_b200 as g
    from gpirt_b200.api import _as_cube, _gpirtMCMC_constant_irf

    n, m, C = 80, 7, 3
    P = make_problem(n, m, C, seed=410, missing=0.1)
    kw = dict(theta_init=P["theta0"], thresholds=P["thr0"], SEED=4, constant_IRF=1, store_f=True, store_fstar=True)
    a = g.gpirtMCMC(P["y"], 4, 1, THIN=2, **kw)[0]
    assert a["theta"].shape == (2, n, 1) and np.all(np.isfinite(a["ll"])) and np.all(a["ll"] < 0)
    b = _gpirtMCMC_constant_irf(_as_cube(P["y"]), 4, 1, 2, 1, None, None, None, None, 1, 10, "Matern", P["theta0"].reshape(n, 1), P["thr0"], 4, True, True, 0, None, 0)[0]
    assert np.array_equal(a["theta"], b["theta"]) and np.array_equal(a["ll"], b["ll"])
    for k in range(2):
        for key in ("beta", "threshold", "f", "fstar"):
            assert np.array_equal(a[key][k], b[key][k]), f"{key} draw {k}"
    plain = g.gpirtMCMC(P["y"], 4, 1, THIN=2, **dict(kw, constant_IRF=0))[0]
    assert not np.array_equal(a["theta"], plain["theta"])  # the f* draw differs, and with it the theta step
