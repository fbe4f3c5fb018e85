"""Golden vectors recorded from the REFERENCE'S OWN CODE (tests/golden/make_golden.py: /root/reference/src/*.cpp compiled
against oracle/refshim, single thread). They travel with the repo, so both the CPU oracle (here, no GPU) and the CUDA path
(-m gpu, on the B200 box where /root/reference does not exist) are checked against the reference's outputs, not only against
each other. The sweep fixtures hold the random streams every unit consumed in the reference run; both sides replay them as
supplied streams.

Tolerances: theta and indexing bit-exact; beta / thresholds / f 1e-9 relative (cos/sin/exp rounding and summation order of
L*z); total log-likelihood 1e-9 relative (north star: 1e-10 on the likelihood itself, tested separately on fixed states below);
f* and the f gathered from it 4e-6 norm-wise per item column (K^-1 f is kappa-limited with kappa ~ 1e10 under the default priors,
SURVEY.md 7 — two correct fp64 codes differ at that level; the oracle itself differs from the reference by 1e-7)."""
import os

import numpy as np
import pytest

from oracle import orc

HERE = os.path.dirname(os.path.abspath(__file__))
SWEEPS = ["sweep_ordinal_c4", "sweep_binary_c2"]


def load(name):
    return dict(np.load(os.path.join(HERE, "golden", name + ".npz")))


def rel(a, b):
    a, b = np.asarray(a), np.asarray(b)
    fin = np.isfinite(b)
    assert np.array_equal(np.isfinite(a), fin)
    return float(np.max(np.abs(a[fin] - b[fin]) / np.maximum(1.0, np.abs(b[fin])))) if fin.any() else 0.0


# ---- CPU: the oracle against the reference's vectors -------------------------------------------------------------
def test_leaves_oracle_matches_reference_vectors():
    G = load("leaves")
    assert np.array_equal(orc.K(G["x1"], G["x2"], G["sds"]), G["K"])
    assert np.array_equal(orc.K_time(G["ts"], G["ts"], 1.3, 2.0, [0.4, 0.2], "Matern"), G["K_time_matern"])
    assert np.array_equal(orc.K_time(G["ts"], G["ts"], 1.3, 2.0, [0.4, 0.2], "RBF"), G["K_time_rbf"])
    assert np.array_equal(orc.pnorm(G["pnorm_x"]), G["pnorm"])
    assert orc.ll_bar_sparse(G["ll_f"], G["ll_y"], G["ll_mu"], G["ll_thr"]) == float(G["ll"])
    assert np.array_equal(orc.threshold_to_delta(G["ll_thr"]), G["delta"])
    assert np.array_equal(orc.delta_to_threshold(G["delta"]), G["thr_back"])
    assert np.array_equal(orc.delta_to_threshold([0.25]), G["thr2"])
    th = G["chol_theta"]
    A = orc.K(th, th, G["chol_sds"][:, 0]) + 1e-6 * np.eye(th.size)
    L, info = orc.chol(A)
    assert info == 0 and np.max(np.abs(L - G["chol_L"])) < 1e-13
    out = np.stack([orc.double_solve(G["chol_L"], G["solve_X"][:, k]) for k in range(3)], axis=1)
    assert rel(out, G["solve_out"]) < 1e-12


def colrel(a, b):
    """norm-wise relative difference per column (item): max |a - b| / max |b|"""
    return float(np.max(np.max(np.abs(a - b), axis=0) / np.maximum(1.0, np.max(np.abs(b), axis=0))))


def replay(G, S, kind):
    """Drives one sampler (oracle or CUDA) through the fixture's sweeps with the recorded streams; yields its state after each."""
    n, m, C = int(G["n"]), int(G["m"]), int(G["C"])
    for it in range(int(G["iters"])):
        zf, uf = G[f"z_f_{it}"], G[f"u_f_{it}"]
        if kind == "orc":
            S.compute_mu()
            assert S.update_cholesky() == 0
            nu = np.asfortranarray(S.L() @ zf)  # rmvnorm_threadsafe: nu_j = L z_j (src/mvnormal.h:17-24)
            S.draw_f(orc.supplied(u=uf), nu=nu)
            assert S.draw_fstar(orc.supplied(z=G[f"z_fstar_{it}"])) == 0
            S.draw_theta(orc.supplied(z=G[f"z_theta_{it}"], u=G[f"u_theta_{it}"]))
            S.gather_f()
            S.draw_beta(orc.supplied(z=G[f"z_beta_{it}"], u=G[f"u_beta_{it}"]))
            S.compute_mu()
            assert S.update_cholesky() == 0
            S.draw_threshold(orc.supplied(z=G[f"z_thr_{it}"], u=G[f"u_thr_{it}"]))
            st = S.state()
            st["thresholds"] = st.pop("thr")
            yield it, st, S.total_ll(), {b: S.diag(b)[0] for b in ("f", "theta", "beta", "thr")}
        else:
            S.draw_f(z=zf, u=uf)
            S.draw_fstar(z=G[f"z_fstar_{it}"])
            S.draw_theta(z=G[f"z_theta_{it}"], u=G[f"u_theta_{it}"])
            S.gather_f()
            S.draw_beta(z=G[f"z_beta_{it}"], u=G[f"u_beta_{it}"])
            S.update_cholesky_cache()
            S.draw_threshold(z=G[f"z_thr_{it}"], u=G[f"u_thr_{it}"])
            _, tot = S.ll_bar()
            yield it, S.get_state(fstar=True), tot, {b: S.diag(b)[0] for b in ("f", "theta", "beta", "thr")}


def check_against_reference(G, it, st, ll, evals, last):
    assert np.array_equal(st["theta"], G["ref_theta"][it]), f"sweep {it}: theta differs from the reference's draw"
    assert rel(st["beta"], G["ref_beta"][:, :, it]) < 1e-9
    assert rel(st["thresholds"], G["ref_threshold"][:, :, it]) < 1e-9
    assert colrel(st["f"], G["ref_f"][:, :, it]) < 4e-6  # f = rows of f* after the gather
    assert abs(ll - G["ref_ll"][it]) <= 1e-7 * abs(G["ref_ll"][it])  # carries the f* difference; the likelihood itself is tested at 1e-10
    # every unit made exactly the number of proposals the reference made (uniforms consumed: u, eps0, one per rejection = 1 + proposals = evals)
    for b in ("f", "theta", "beta", "thr"):
        nu = G[f"nu_{b}_{it}"]
        assert np.array_equal(np.where(nu > 0, evals[b], 0), nu), f"sweep {it}: block {b} made a different number of proposals"
    if last:
        assert colrel(st["fstar"], G["ref_fstar_last"]) < 4e-6


@pytest.mark.parametrize("name", SWEEPS)
def test_oracle_replays_reference_sweeps(name):
    G = load(name)
    O = orc.Sampler(G["y"], int(G["C"]), G["beta_prior_sds"], G["theta_prior_sds"])
    O.set_state(theta=np.clip(G["theta_in"], -5, 5), beta=G["beta0"], f=G["f0"], thr=G["thr0"], fstar=G["fstar0"])
    for it, st, ll, ev in replay(G, O, "orc"):
        check_against_reference(G, it, st, ll, ev, it == int(G["iters"]) - 1)


# ---- GPU: the CUDA path against the same vectors -------------------------------------------------------------------
@pytest.mark.gpu
def test_leaves_cuda_matches_reference_vectors():
    import gpirt_b200 as g

    G = load("leaves")
    assert rel(g.K(G["x1"], G["x2"], G["sds"]), G["K"]) < 1e-14
    assert rel(g.K_time(G["ts"], G["ts"], 1.3, 2.0, [0.4, 0.2], "Matern"), G["K_time_matern"]) < 1e-14
    assert rel(g.K_time(G["ts"], G["ts"], 1.3, 2.0, [0.4, 0.2], "RBF"), G["K_time_rbf"]) < 1e-14
    from gpirt_b200.api import debug_probit

    ph, _ = debug_probit(G["pnorm_x"])
    assert np.max(np.abs(ph - G["pnorm"])) < 1e-15  # absolute: the likelihood adds 1e-6 inside the log
    assert np.array_equal(g.threshold_to_delta(G["ll_thr"]), G["delta"])
    assert np.array_equal(g.delta_to_threshold(G["delta"]), G["thr_back"])
    th = G["chol_theta"]
    A = g.K(th, th, G["chol_sds"][:, 0]) + 1e-6 * np.eye(th.size)
    assert np.max(np.abs(g.chol(A) - G["chol_L"])) < 1e-10  # well-conditioned fixture: element-wise 1e-10
    assert rel(g.double_solve(G["chol_L"], G["solve_X"]), G["solve_out"]) < 1e-10
    # ll_bar_sparse through a one-item sampler
    n, C = G["ll_f"].size, G["ll_thr"].size - 1
    S = g.Sampler(n, 1, C)
    S.set_data(G["ll_y"].reshape(n, 1))
    S.set_priors(np.ones((3, 1)), np.zeros((2, n)))
    # mu = X beta with theta = mu, beta = (0, 1, 0) reproduces the fixture's mu exactly
    S.set_state(theta=np.clip(G["ll_mu"], -5, 5), beta=np.array([[0.0], [1.0], [0.0]]), f=G["ll_f"].reshape(n, 1), thresholds=G["ll_thr"].reshape(1, -1))
    _, tot = S.ll_bar()
    assert abs(tot - float(G["ll"])) <= 1e-10 * abs(float(G["ll"]))


@pytest.mark.gpu
@pytest.mark.parametrize("name", SWEEPS)
def test_cuda_replays_reference_sweeps(name):
    import gpirt_b200 as g

    G = load(name)
    n, m, C = int(G["n"]), int(G["m"]), int(G["C"])
    S = g.Sampler(n, m, C)
    S.set_data(G["y"])
    S.set_priors(G["beta_prior_sds"], G["theta_prior_sds"])
    S.set_state(theta=G["theta_in"], beta=G["beta0"], f=G["f0"], thresholds=G["thr0"], fstar=G["fstar0"])  # set_state clamps theta (src/gpirtMCMC.cpp:133)
    for it, st, ll, ev in replay(G, S, "cuda"):
        check_against_reference(G, it, st, ll, ev, it == int(G["iters"]) - 1)


# ---- horizon > 1: the reference's own H = 3 runs (RDM and GP-over-time regimes), streams recorded per session --------------------
PANELS = ["panel_rdm_h3", "panel_gp_h3"]


def replay_panel(G, P, kind):
    """P: oracle Panel (kind 'orc') or gpirt_b200.Panel (kind 'cuda'); yields (sweep, state, ll, theta evals) after each replayed sweep."""
    H, rdm = int(G["H"]), float(G["ls"]) <= 0.1
    for it in range(int(G["iters"])):
        k = lambda blk, h: (G[f"z_{blk}_{it}_{h}"], G.get(f"u_{blk}_{it}_{h}"))
        if kind == "orc":
            for S in P.S:
                S.compute_mu()
                assert S.update_cholesky() == 0
            for h, S in enumerate(P.S):
                z, u = k("f", h)
                S.draw_f(orc.supplied(u=u), nu=np.asfortranarray(S.L() @ z))
            for h, S in enumerate(P.S):
                assert S.draw_fstar(orc.supplied(z=k("fstar", h)[0])) == 0
            if rdm:
                P.draw_theta([orc.supplied(z=k("theta", h)[0], u=k("theta", h)[1]) for h in range(H)])
                ev = np.stack([S.diag("theta")[0] for S in P.S])
            else:
                P.draw_theta(orc.supplied(z=G[f"z_theta_{it}"], u=G[f"u_theta_{it}"]))
                ev = P.S[0].diag("theta")[0]
            for S in P.S:
                S.gather_f()
            for h, S in enumerate(P.S):
                S.draw_beta(orc.supplied(z=k("beta", h)[0], u=k("beta", h)[1]))
            for S in P.S:
                S.compute_mu()
                assert S.update_cholesky() == 0
            for h, S in enumerate(P.S):
                S.draw_threshold(orc.supplied(z=k("thr", h)[0], u=k("thr", h)[1]))
            st = P.state()
            st["thresholds"] = st.pop("thr")
            yield it, st, P.total_ll(), ev
        else:
            for h, S in enumerate(P.S):
                z, u = k("f", h)
                S.draw_f(z=z, u=u)
            for h, S in enumerate(P.S):
                S.draw_fstar(z=k("fstar", h)[0])
            if rdm:
                P.draw_theta(z=[k("theta", h)[0] for h in range(H)], u=[k("theta", h)[1] for h in range(H)])
            else:
                P.draw_theta(z=G[f"z_theta_{it}"], u=G[f"u_theta_{it}"])
            ev = P.diag_theta()[0]
            for S in P.S:
                S.gather_f()
            for h, S in enumerate(P.S):
                S.draw_beta(z=k("beta", h)[0], u=k("beta", h)[1])
            for S in P.S:
                S.update_cholesky_cache()
            for h, S in enumerate(P.S):
                S.draw_threshold(z=k("thr", h)[0], u=k("thr", h)[1])
            ll = sum(S.ll_bar()[1] for S in P.S)
            yield it, P.get_state(fstar=True), ll, ev


def check_panel(G, it, st, ll, ev, last):
    H, rdm = int(G["H"]), float(G["ls"]) <= 0.1
    assert np.array_equal(st["theta"], G["ref_theta"][it]), f"sweep {it}: theta differs from the reference's draw"
    assert rel(st["beta"], G["ref_beta"][..., it]) < 1e-9 and rel(st["thresholds"], G["ref_threshold"][..., it]) < 1e-9
    for h in range(H):
        assert colrel(st["f"][:, :, h], G["ref_f"][:, :, h, it]) < 4e-6
        if last:
            assert colrel(st["fstar"][:, :, h], G["ref_fstar_last"][:, :, h]) < 4e-6
    assert abs(ll - G["ref_ll"][it]) <= 1e-7 * abs(G["ref_ll"][it])
    if rdm:
        nu = np.stack([G[f"nu_theta_{it}_{h}"] for h in range(H)])
    else:
        nu = G[f"nu_theta_{it}"]
    assert np.array_equal(np.where(nu > 0, ev, 0), nu), f"sweep {it}: the theta step made a different number of proposals"


@pytest.mark.parametrize("name", PANELS)
def test_oracle_panel_replays_reference_horizon_3_sweeps(name):
    G = load(name)
    O = orc.Panel(G["y"], int(G["C"]), G["beta_prior_sds"], G["theta_prior_sds"], theta_os=float(G["os"]), theta_ls=float(G["ls"]), KERNEL=str(G["kernel"]))
    O.set_state(theta=np.clip(G["theta_in"], -5, 5), beta=G["beta0"], f=G["f0"], thr=G["thr0"], fstar=G["fstar0"])
    for it, st, ll, ev in replay_panel(G, O, "orc"):
        check_panel(G, it, st, ll, ev, it == int(G["iters"]) - 1)


@pytest.mark.gpu
@pytest.mark.parametrize("name", PANELS)
def test_cuda_panel_replays_reference_horizon_3_sweeps(name):
    import gpirt_b200 as g

    G = load(name)
    n, m, C, H = int(G["n"]), int(G["m"]), int(G["C"]), int(G["H"])
    P = g.Panel(n, m, C, H, theta_os=float(G["os"]), theta_ls=float(G["ls"]), KERNEL=str(G["kernel"]))
    P.set_data(G["y"])
    P.set_priors(G["beta_prior_sds"], G["theta_prior_sds"])
    P.set_state(theta=G["theta_in"], beta=G["beta0"], f=G["f0"], thresholds=G["thr0"], fstar=G["fstar0"])
    for it, st, ll, ev in replay_panel(G, P, "cuda"):
        check_panel(G, it, st, ll, ev, it == int(G["iters"]) - 1)


# ---- constant_IRF == 1: the reference's own H = 2 runs, streams recorded for the stacked blocks / per session -----------------------------
CSTS = ["cst_rdm_h2", "cst_gp_h2", "cst_cst_h1"]  # the last: one session, R's default theta_ls = 10 (CST regime)


@pytest.mark.parametrize("name", CSTS)
def test_oracle_constant_irf_panel_replays_reference_sweeps(name):
    """The oracle's ConstantPanel (oracle/orc.py) driven by the streams the reference consumed (tests/golden/make_golden.py cst_case): theta
    bit-exact, identical proposal counts in the stacked f / threshold blocks, beta / thresholds 1e-9. The same fixtures are
    replayed by the CUDA composition in test_cuda_constant_irf_panel_replays_reference_sweeps."""
    G = load(name)
    n, m, C, H = int(G["n"]), int(G["m"]), int(G["C"]), int(G["H"])
    rdm = float(G["ls"]) <= 0.1 or H == 1  # per-session theta streams
    O = orc.ConstantPanel(G["y"], C, G["beta_prior_sds"], G["theta_prior_sds"], theta_os=1.0, theta_ls=float(G["ls"]), KERNEL="Matern")
    O.init_constant(np.clip(G["theta_in"], -5, 5), G["beta0"], G["f0"], G["thr0"])
    for S in O.S:
        S.set_state(fstar=G["fstar0"])
    for it in range(int(G["iters"])):
        O.draw_f(orc.supplied(z=G[f"z_f_{it}"], u=G[f"u_f_{it}"]))
        ev_f = O.T.diag("f")[0]
        assert O.draw_fstar(orc.supplied(z=G[f"z_fstar_{it}"])) == 0
        if rdm:
            O.draw_theta([orc.supplied(z=G[f"z_theta_{it}_{h}"], u=G[f"u_theta_{it}_{h}"]) for h in range(H)])
        else:
            O.draw_theta(orc.supplied(z=G[f"z_theta_{it}"], u=G[f"u_theta_{it}"]))
        for S in O.S:
            S.gather_f()
        for h, S in enumerate(O.S):
            S.draw_beta(orc.supplied(z=G[f"z_beta_{it}_{h}"], u=G[f"u_beta_{it}_{h}"]))
        for S in O.S:
            S.compute_mu()
        O.draw_threshold(orc.supplied(z=G[f"z_thr_{it}"], u=G[f"u_thr_{it}"]))
        ev_thr = O.T.diag("thr")[0]
        s = O.state()
        assert np.array_equal(s["theta"], G["ref_theta"][it]), f"sweep {it}: theta differs from the reference's draw"
        assert rel(s["beta"], G["ref_beta"][..., it]) < 1e-9 and rel(s["thr"], G["ref_threshold"][..., it]) < 1e-9
        assert abs(O.total_ll() - G["ref_ll"][it]) <= 1e-7 * abs(G["ref_ll"][it])
        for ev, key in ((ev_f, "f"), (ev_thr, "thr")):
            nu = G[f"nu_{key}_{it}"]
            assert np.array_equal(np.where(nu > 0, ev, 0), nu), f"sweep {it}: stacked block {key} made a different number of proposals"
        assert np.max(np.abs(s["f"] - G["ref_f"][..., it])) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("name", CSTS)
def test_cuda_constant_irf_panel_replays_reference_sweeps(name):
    """gpirt_b200.ConstantPanel (panel + stacked sampler with per-row means + 100-respondent inducing sampler + gpirt_interp1) driven by the streams
    the reference consumed: theta bit-exact, identical proposal counts in the stacked f / threshold blocks, beta / thresholds 1e-9; f* is the
    solution of the 100-point inducing system (kappa ~ 1e12): compared to the reference's at 5e-3 absolute."""
    from gpirt_b200.constant_panel import ConstantPanel

    G = load(name)
    n, m, C, H = int(G["n"]), int(G["m"]), int(G["C"]), int(G["H"])
    rdm = float(G["ls"]) <= 0.1 or H == 1  # per-session theta streams
    P = ConstantPanel(n, m, C, H, theta_os=1.0, theta_ls=float(G["ls"]), KERNEL="Matern")
    P.set_data(G["y"])
    P.set_priors(G["beta_prior_sds"], G["theta_prior_sds"])
    P.init_constant(G["theta_in"], G["beta0"], G["f0"], G["thr0"])
    for S in P.S:
        S.set_state(fstar=G["fstar0"])
    for it in range(int(G["iters"])):
        P.draw_f(z=G[f"z_f_{it}"], u=G[f"u_f_{it}"])
        ev_f = P.T.diag("f")[0]
        P.draw_fstar(z=G[f"z_fstar_{it}"])
        if rdm:
            P.draw_theta(z=[G[f"z_theta_{it}_{h}"] for h in range(H)], u=[G[f"u_theta_{it}_{h}"] for h in range(H)])
        else:
            P.draw_theta(z=G[f"z_theta_{it}"], u=G[f"u_theta_{it}"])
        P.gather_f()
        P.draw_beta(z=[G[f"z_beta_{it}_{h}"] for h in range(H)], u=[G[f"u_beta_{it}_{h}"] for h in range(H)])
        for S in P.S:
            S.update_cholesky_cache()
        P.draw_threshold(z=G[f"z_thr_{it}"], u=G[f"u_thr_{it}"])
        ev_thr = P.T.diag("thr")[0]
        s = P.get_state(fstar=True)
        assert np.array_equal(s["theta"], G["ref_theta"][it]), f"sweep {it}: theta differs from the reference's draw"
        assert rel(s["beta"], G["ref_beta"][..., it]) < 1e-9 and rel(s["thresholds"], G["ref_threshold"][..., it]) < 1e-9
        assert abs(P.total_ll() - G["ref_ll"][it]) <= 2e-5 * abs(G["ref_ll"][it])  # carries the f* difference of the kappa ~ 1e12 inducing system
        for ev, key in ((ev_f, "f"), (ev_thr, "thr")):
            nu = G[f"nu_{key}_{it}"]
            assert np.array_equal(np.where(nu > 0, ev, 0), nu), f"sweep {it}: stacked block {key} made a different number of proposals"
        assert np.max(np.abs(s["f"] - G["ref_f"][..., it])) < 5e-3
        assert np.array_equal(s["fstar"][:, :, 0], s["fstar"][:, :, -1])
    assert np.max(np.abs(s["fstar"] - G["ref_fstar_last"])) < 5e-3
    P.close()


@pytest.mark.gpu
def test_interp1_on_the_gpu_matches_the_oracle_restatement():
    import gpirt_b200.api as api

    rng = np.random.default_rng(12)
    X = np.round(rng.standard_normal(400) * 1.5, 2)  # duplicated abscissae, as theta on the 0.01 grid
    Y = np.asfortranarray(rng.standard_normal((400, 7)))
    XI = np.concatenate([np.linspace(X.min(), X.max(), 100), [X.min() - 1.0, X.max() + 0.5, X[17]]])
    got = api.interp1(X, Y, XI)
    order = np.argsort(X, kind="stable")
    xs, first = np.unique(X[order], return_index=True)  # first of each run in stable order = lowest row index
    ys = Y[order[first]]
    want = np.full_like(got, np.nan)
    for q, x in enumerate(XI):
        if x < xs[0] or x > xs[-1]:
            continue
        hi = int(np.searchsorted(xs, x, side="left"))
        if xs[hi] == x:
            want[q] = ys[hi]
        else:
            w = (x - xs[hi - 1]) / (xs[hi] - xs[hi - 1])
            want[q] = (1.0 - w) * ys[hi - 1] + w * ys[hi]
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(got[~np.isnan(want)], want[~np.isnan(want)])
