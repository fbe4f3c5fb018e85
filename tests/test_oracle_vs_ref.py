"""Pins the CPU oracle to the REFERENCE'S OWN CODE: oracle/_ref/libgpirt_ref.so is /root/reference/src/*.cpp compiled where it
lies against the stand-in headers of oracle/refshim (oracle/Makefile `ref`). Runs wherever that library exists (this container;
it also travels to the GPU box as a built file); skipped otherwise — tests/test_golden.py then still checks the recorded vectors.

Single OpenMP thread on both sides: the reference's RNG streams are keyed by thread id (src/gpirt.h:178-182)."""
import os

import numpy as np
import pytest

from oracle import orc, ref
from tests.problems import make_problem

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")


@pytest.fixture(autouse=True)
def one_thread():
    orc.set_threads(1)
    ref.set_threads(1)
    yield
    orc.set_threads(os.cpu_count() or 1)


def test_leaves_bit_identical():
    rng = np.random.default_rng(3)
    x1, x2, sds = rng.standard_normal(33) * 2.5, rng.standard_normal(21) * 2.5, [3.0, 3.0, 3.0]
    assert np.array_equal(orc.K(x1, x2, sds), ref.K(x1, x2, sds))
    ts = np.arange(6.0)
    for kern in ("Matern", "RBF", "other"):
        assert np.array_equal(orc.K_time(ts, ts, 0.9, 1.7, [0.2, 0.3], kern), ref.K_time(ts, ts, 0.9, 1.7, [0.2, 0.3], kern))
    x = np.concatenate([np.linspace(-39, 9, 3000), rng.standard_normal(1000)])
    assert np.array_equal(orc.pnorm(x), ref.pnorm(x))
    for C in (2, 3, 7):
        d = rng.standard_normal(C - 1)
        t = ref.delta_to_threshold(d)
        assert np.array_equal(orc.delta_to_threshold(d), t)
        assert np.array_equal(orc.threshold_to_delta(t), ref.threshold_to_delta(t))
    n = 200
    f, mu = rng.standard_normal(n) * 2, rng.standard_normal(n)
    y = rng.integers(1, 6, n).astype(float)
    y[rng.random(n) < 0.1] = np.nan
    thr = np.array([-np.inf, -1.0, -0.3, 0.4, 1.2, np.inf])
    assert orc.ll_bar_sparse(f, y, mu, thr) == ref.ll_bar_sparse(f, y, mu, thr)
    assert ref.ll_bar_sparse(f, np.full(n, np.nan), mu, thr) == 0.0  # an item nobody answered


def test_cholesky_cache_and_double_solve():
    rng = np.random.default_rng(4)
    n = 90
    theta = np.clip(rng.standard_normal(n), -5, 5)
    sds = np.full((3, 4), 3.0)
    L = ref.update_cholesky_cache(theta, sds)[0][:, :, 0]
    Lo, info = orc.chol(orc.K(theta, theta, sds[:, 0]) + 1e-6 * np.eye(n))
    assert info == 0 and np.array_equal(L, Lo)  # same unblocked column algorithm on both sides (LAPACK in the real reference)
    x = rng.standard_normal(n)
    assert np.array_equal(orc.double_solve(L, x), ref.double_solve(L, x)[:, 0])
    with pytest.raises(RuntimeError, match="chol"):  # arma::chol throws on a non-PD matrix -> R error
        ref.update_cholesky_cache(np.array([0.0, 0.0]), np.array([[0.0], [0.0], [np.nan]]))


@pytest.mark.parametrize("C,seed", [(2, 1), (5, 2), (12, 3), (21, 4)])  # 21 = GPIRT_MAX_C
def test_blocks_same_state_and_same_rng_consumption(C, seed):
    n, m = 70, 9
    P = make_problem(n, m, C, seed=seed, missing=0.15, on_grid=True)
    rng = P["rng"]
    y, bps, tps = P["y"], P["beta_prior_sds"], np.asfortranarray(rng.random((2, n)))
    theta, beta, thr = P["theta0"], P["beta0"], P["thr0"]
    f = np.asfortranarray(rng.standard_normal((n, m)))
    fstar = np.asfortranarray(rng.standard_normal((orc.NGRID, m)))

    def fresh():
        O = orc.Sampler(y, C, bps, tps)
        O.set_state(theta=theta, beta=beta, f=f, thr=thr, fstar=fstar)
        O.compute_mu()
        assert O.update_cholesky() == 0
        return O

    mu, mu_star = fresh().mu()
    O = fresh(); O.mt_seed_all(777); O.draw_f(orc.mt())
    assert np.array_equal(O.state()["f"], ref.draw_f(777, f, theta, y, bps, mu, thr)[:, :, 0])
    O = fresh(); O.mt_seed_all(778); O.draw_theta(orc.mt())
    assert np.array_equal(O.state()["theta"], ref.draw_theta(778, y, theta, tps, fstar, mu_star, thr, bps)[:, 0])
    O = fresh(); O.mt_seed_all(779); O.draw_beta(orc.mt())
    assert np.array_equal(O.state()["beta"], ref.draw_beta(779, beta, theta, y, f, np.zeros((3, m)), bps, thr)[:, :, 0])
    O = fresh(); O.mt_seed_all(780); O.draw_threshold(orc.mt())
    assert np.array_equal(O.state()["thr"], ref.draw_threshold(780, thr, y, f, mu)[:, :, 0])
    O = fresh(); O.mt_seed_all(781); assert O.draw_fstar(orc.mt()) == 0
    R = ref.draw_fstar(781, f, theta, bps, mu_star)[:, :, 0]
    assert np.max(np.abs(O.state()["fstar"] - R)) < 1e-5  # K^-1 f is kappa-limited (kappa ~ 1e10 with sds = 3)


def test_gpirtMCMC_same_draws_over_several_sweeps():
    n, m, C, iters = 64, 10, 5, 5
    P = make_problem(n, m, C, seed=9, missing=0.1)
    rng = np.random.default_rng(10)
    z = rng.standard_normal(3 * m + n * m)
    bpm, bps, tpm, tps = np.zeros((3, m)), P["beta_prior_sds"], np.zeros((2, n)), P["theta_prior_sds"]
    R = ref.gpirtMCMC(P["y"], P["theta0"], iters, 1, 1, bpm, bps, tpm, tps, 1.0, 10.0, "Matern", P["thr0"], 0, z, store_f=True)
    theta0 = np.clip(P["theta0"], -5, 5)
    beta0 = bpm + bps * z[:3 * m].reshape(m, 3).T
    L0 = ref.update_cholesky_cache(theta0, bps)[0][:, :, 0]
    f0 = np.stack([L0 @ z[3 * m + j * n:3 * m + (j + 1) * n] for j in range(m)], axis=1)
    O = orc.Sampler(P["y"], C, bps, tps)
    O.set_state(theta=theta0, beta=beta0, f=f0, thr=P["thr0"])
    O.compute_mu()
    assert O.update_cholesky() == 0
    O.draw_fstar(orc.mt())  # initial draw on the unseeded generators (src/gpirtMCMC.cpp:227)
    info, th, ll = O.run_mt(0, iters + 1)
    assert info == 0
    assert np.array_equal(th[:, 1:].T, R["theta"][:, :, 0])  # burn-in sweep dropped by the reference (src/gpirtMCMC.cpp:334)
    assert np.allclose(ll[1:], R["ll"], rtol=1e-9, atol=0)
    s = O.state()
    assert np.array_equal(s["beta"], R["beta"][:, :, 0, -1])
    assert np.array_equal(s["thr"][:, 1:-1], R["threshold"][:, 1:-1, 0, -1])
    assert np.max(np.abs(s["f"] - R["f"][:, :, 0, -1])) < 1e-5


def test_gpirtMCMC_with_a_respondent_and_an_item_without_answers():
    """Ragged input: respondent 5 answered nothing (empty obs_items: every theta proposal has ll = 0 and the first one is accepted,
    src/draw-theta.cpp:4-24,44-84), item 2 was answered by nobody (draw_beta skips it, src/draw-beta.cpp:73), item 4 by one respondent only."""
    n, m, C, iters = 40, 7, 3, 4
    P = make_problem(n, m, C, seed=19, missing=0.3)
    P["y"][5, :] = np.nan
    P["y"][:, 2] = np.nan
    P["y"][:, 4] = np.nan
    P["y"][11, 4] = 2.0
    z = np.random.default_rng(20).standard_normal(3 * m + n * m)
    bpm, bps, tpm, tps = np.zeros((3, m)), P["beta_prior_sds"], np.zeros((2, n)), P["theta_prior_sds"]
    R = ref.gpirtMCMC(P["y"], P["theta0"], iters, 0, 1, bpm, bps, tpm, tps, 1.0, 10.0, "Matern", P["thr0"], 0, z)
    theta0 = np.clip(P["theta0"], -5, 5)
    beta0 = bpm + bps * z[:3 * m].reshape(m, 3).T
    L0 = ref.update_cholesky_cache(theta0, bps)[0][:, :, 0]
    f0 = np.stack([L0 @ z[3 * m + j * n:3 * m + (j + 1) * n] for j in range(m)], axis=1)
    O = orc.Sampler(P["y"], C, bps, tps)
    O.set_state(theta=theta0, beta=beta0, f=f0, thr=P["thr0"])
    O.compute_mu()
    assert O.update_cholesky() == 0
    O.draw_fstar(orc.mt())
    info, th, ll = O.run_mt(0, iters)
    assert info == 0
    assert np.array_equal(th.T, R["theta"][:, :, 0])
    assert np.allclose(ll, R["ll"], rtol=1e-9, atol=0)
    s = O.state()
    assert np.array_equal(s["beta"], R["beta"][:, :, 0, -1]) and np.array_equal(s["beta"][:, 2], beta0[:, 2])
    assert np.array_equal(s["thr"][:, 1:-1], R["threshold"][:, 1:-1, 0, -1])


def test_recover_fstar():
    n, m, C = 50, 6, 3
    P = make_problem(n, m, C, seed=12, missing=0.1, on_grid=True)
    rng = P["rng"]
    f = np.asfortranarray(rng.standard_normal((n, m)))
    beta2 = np.asfortranarray(rng.standard_normal((2, m)))
    bps3 = np.vstack([np.full((2, m), 0.5), np.zeros((1, m))])  # third row: what K() reads past the 2-row matrix is taken as 0
    R = ref.recover_fstar(321, f, P["y"], P["theta0"], beta2, P["thr0"], np.zeros((3, m)), bps3)[:, :, 0]
    O = orc.Sampler(P["y"], C, bps3, None, p_beta=2)
    O.set_state(theta=P["theta0"], beta=beta2, f=f, thr=P["thr0"])
    O.compute_mu()
    assert O.update_cholesky() == 0
    O.mt_seed_all(321)
    O.draw_f(orc.mt())
    assert O.draw_fstar(orc.mt()) == 0
    assert np.max(np.abs(O.state()["fstar"] - R)) < 1e-6


def _recover_panel_inputs(n, m, C, H):
    Ps = [make_problem(n, m, C, seed=40 + h, missing=0.1, on_grid=True) for h in range(H)]
    rng = Ps[0]["rng"]
    y = np.stack([P["y"] for P in Ps], axis=2)
    theta = np.stack([P["theta0"] for P in Ps], axis=1)
    thr = np.stack([P["thr0"] for P in Ps], axis=2)
    f = np.asfortranarray(rng.standard_normal((n, m, H)))
    beta2 = np.asfortranarray(rng.standard_normal((2, m, H)))
    bps3 = np.vstack([np.full((2, m), 0.5), np.zeros((1, m))])
    beta3 = np.concatenate([beta2, np.zeros((1, m, H))], axis=0)  # a zero quadratic row: the same mu and mu* as the 2-row beta
    return y, theta, thr, f, beta2, beta3, bps3


def test_recover_fstar_over_a_panel():
    """recover_fstar at horizon 3, constant_IRF 0 (src/recover_fstar.cpp:86-90 -> src/draw-f.cpp:141-179, src/draw-fstar.cpp:127-168): sessions in order."""
    n, m, C, H = 40, 5, 3, 3
    y, theta, thr, f, beta2, beta3, bps3 = _recover_panel_inputs(n, m, C, H)
    R = ref.recover_fstar(77, f, y, theta, beta2, thr, np.zeros((3, m)), bps3)
    O = orc.Panel(y, C, bps3, None, theta_ls=1.0)
    O.set_state(theta=theta, beta=beta3, f=f, thr=thr)
    O.compute_mu()
    assert O.update_cholesky() == 0
    O.mt_seed_all(77)
    for S in O.S:
        S.draw_f(orc.mt())
    for S in O.S:
        assert S.draw_fstar(orc.mt()) == 0
    assert np.max(np.abs(O.state()["fstar"] - R)) < 1e-6


@pytest.mark.parametrize("H", [2, 1])
def test_recover_fstar_constant_IRF(H):
    """recover_fstar at horizon 2 (and with one session), constant_IRF 1: stacked draw_f with mu_h = [1, theta_h] beta_h (src/draw-f.cpp:84-139), inducing-point f*
    with mu* of session 0, copied to every slice (src/draw-fstar.cpp:58-125)."""
    n, m, C = 40, 5, 3
    y, theta, thr, f, beta2, beta3, bps3 = _recover_panel_inputs(n, m, C, H)
    R = ref.recover_fstar(78, f, y, theta, beta2, thr, np.zeros((3, m)), bps3, constant_IRF=1)
    assert np.array_equal(R[:, :, 0], R[:, :, -1])
    O = orc.ConstantPanel(y, C, bps3, None, theta_ls=1.0)
    O.set_state(theta=theta, beta=beta3, f=f, thr=thr)
    O.compute_mu()
    O.mt_seed_all(78)
    O.draw_f(orc.mt())
    assert O.draw_fstar(orc.mt()) == 0
    got = O.state()["fstar"]
    assert np.array_equal(got[:, :, 0], got[:, :, -1])
    assert np.max(np.abs(got - R)) < 2e-5  # the inducing-point solve runs at kappa ~ 1e12


@pytest.mark.parametrize("regime,ls,kernel", [("RDM", 0.05, "Matern"), ("GP", 2.0, "Matern"), ("GP", 4.0, "RBF")])
def test_gpirtMCMC_horizon_3_same_draws(regime, ls, kernel):
    """horizon > 1 (dynamic GPIRT): the oracle's panel of single-session samplers + joint theta step against the reference's own
    gpirtMCMC run, one thread: theta / beta / thresholds bit for bit in the RDM (src/draw-theta.cpp:171-205) and GP-over-time
    (:206-235) regimes."""
    n, m, C, H, iters = 40, 7, 4, 3, 4
    rng = np.random.default_rng(5)
    Ps = [make_problem(n, m, C, seed=20 + h, missing=0.15) for h in range(H)]
    y = np.stack([P["y"] for P in Ps], axis=2)
    theta_in = np.stack([P["theta0"] for P in Ps], axis=1)
    thr0 = np.stack([Ps[0]["thr0"]] * H, axis=2)
    z = rng.standard_normal(H * (3 * m + n * m))
    bpm, bps, tpm = np.zeros((3, m)), Ps[0]["beta_prior_sds"], np.zeros((2, n))
    tps = np.asfortranarray(np.vstack([np.full((1, n), 0.4), np.zeros((1, n))]))
    os_ = 1.3
    R = ref.gpirtMCMC(y, theta_in, iters, 0, 1, bpm, bps, tpm, tps, os_, ls, kernel, thr0, 0, z, store_f=True, store_fstar=True)
    # initial state as the reference draws it (src/gpirtMCMC.cpp:149-163): beta for every (h, j, p) first, then f = L_h z per (h, j)
    theta0 = np.clip(theta_in, -5, 5)
    beta0 = np.stack([bpm + bps * z[h * 3 * m:(h + 1) * 3 * m].reshape(m, 3).T for h in range(H)], axis=2)
    zf = z[H * 3 * m:]
    L0 = ref.update_cholesky_cache(theta0, bps)[0]
    f0 = np.stack([np.stack([L0[:, :, h] @ zf[(h * m + j) * n:(h * m + j + 1) * n] for j in range(m)], axis=1) for h in range(H)], axis=2)
    O = orc.Panel(y, C, bps, tps, theta_os=os_, theta_ls=ls, KERNEL=kernel)
    assert O.regime == (1 if regime == "RDM" else 2)
    Lt = ref.update_cholesky_cache(theta0, bps, os_, ls, kernel)[1][:, :, 0]
    if regime == "GP":  # L_time is only formed when 0.1 < ls < 3H (src/cholesky-cache.cpp:27)
        assert np.array_equal(np.tril(O.L_time), np.tril(Lt)) and np.all(np.triu(O.L_time, 1) == 0)
    O.set_state(theta=theta0, beta=beta0, f=f0, thr=thr0)
    O.compute_mu()
    assert O.update_cholesky() == 0
    for S in O.S:
        assert S.draw_fstar(orc.mt()) == 0  # initial draw_fstar on the unseeded generators (src/gpirtMCMC.cpp:227)
    for it in range(iters):
        O.mt_seed_all(it * 10000)
        assert O.sweep(lambda block, h: orc.mt()) == 0
        s = O.state()
        assert np.array_equal(s["theta"], R["theta"][it]), f"sweep {it}"
        assert np.array_equal(s["beta"], R["beta"][..., it])
        assert np.array_equal(s["thr"][:, 1:-1], R["threshold"][:, 1:-1, :, it])
        assert abs(O.total_ll() - R["ll"][it]) <= 1e-9 * abs(R["ll"][it])
    assert np.max(np.abs(s["f"] - R["f"][..., -1])) < 1e-5
    assert np.max(np.abs(s["fstar"] - R["fstar"][..., -1])) < 1e-5


@pytest.mark.parametrize("ls,os_,tp", [(10.0, 1.0, 0.4), (0.05, 1.0, 0.4), (2.5, 1.3, 0.4), (2.5, 1.0, 0.0)])
def test_gpirtMCMC_one_session_in_every_theta_regime(ls, os_, tp):
    """horizon 1 with theta_ls in the CST (>= 3), RDM (<= 0.1) and GP (between) branches of src/draw-theta.cpp:112,171,206 and non-zero
    theta_prior_sds: the oracle's single-session sampler against the reference's own run, theta bit for bit."""
    n, m, C, iters = 64, 10, 5, 4
    P = make_problem(n, m, C, seed=9, missing=0.1)
    z = np.random.default_rng(10).standard_normal(3 * m + n * m)
    tps = np.asfortranarray(np.vstack([np.full((1, n), tp), np.zeros((1, n))]))
    bpm, bps, tpm = np.zeros((3, m)), P["beta_prior_sds"], np.zeros((2, n))
    R = ref.gpirtMCMC(P["y"], P["theta0"], iters, 1, 1, bpm, bps, tpm, tps, os_, ls, "Matern", P["thr0"], 0, z)
    theta0 = np.clip(P["theta0"], -5, 5)
    beta0 = bpm + bps * z[:3 * m].reshape(m, 3).T
    L0 = ref.update_cholesky_cache(theta0, bps)[0][:, :, 0]
    f0 = np.stack([L0 @ z[3 * m + j * n:3 * m + (j + 1) * n] for j in range(m)], axis=1)
    O = orc.Sampler(P["y"], C, bps, tps, theta_regime=1 if 0.1 < ls < 3.0 else 0, theta_os=os_)
    O.set_state(theta=theta0, beta=beta0, f=f0, thr=P["thr0"])
    O.compute_mu()
    assert O.update_cholesky() == 0
    O.draw_fstar(orc.mt())
    info, th, ll = O.run_mt(0, iters + 1)
    assert info == 0
    assert np.array_equal(th[:, 1:].T, R["theta"][:, :, 0])
    assert np.allclose(ll[1:], R["ll"], rtol=1e-9, atol=0)


@pytest.mark.parametrize("regime,ls,H", [("RDM", 0.05, 2), ("GP", 2.5, 2), ("CST", 10.0, 1), ("GP", 2.5, 1)])
def test_gpirtMCMC_constant_IRF_same_draws(regime, ls, H):
    """constant_IRF == 1 (src/draw-f.cpp:84-139, src/draw-fstar.cpp:58-125, src/draw_threshold.cpp:5-96,181-205, init src/gpirtMCMC.cpp:164-200):
    the oracle's ConstantPanel (panel + one stacked sampler over the n*H rows + the inducing-point f* draw) against the reference's own run,
    one thread: theta / beta / thresholds bit for bit; f* is the solution of a 100-point inducing system with kappa ~ 1e12 (1e-3 absolute)."""
    n, m, C, iters = 30, 5, 4, 3
    Ps = [make_problem(n, m, C, seed=70 + h, missing=0.1) for h in range(H)]
    y = np.asfortranarray(np.stack([P["y"] for P in Ps], axis=2))
    theta_in = np.asfortranarray(np.stack([P["theta0"] for P in Ps], axis=1))
    thr0 = np.asfortranarray(np.stack([Ps[0]["thr0"]] * H, axis=2))
    z = np.random.default_rng(5).standard_normal(3 * m + n * m)  # the constant branch draws beta for session 0 and ONE f0 per item (:176-188)
    bpm, bps, tpm, tps = np.zeros((3, m)), Ps[0]["beta_prior_sds"], np.zeros((2, n)), np.zeros((2, n))
    R = ref.gpirtMCMC(y, theta_in, iters, 0, 1, bpm, bps, tpm, tps, 1.0, ls, "Matern", thr0, 1, z, store_f=True, store_fstar=True)
    theta0 = np.clip(theta_in, -5, 5)
    beta0 = np.asfortranarray(bpm + bps * z[:3 * m].reshape(m, 3).T)
    L0 = ref.update_cholesky_cache(theta0, bps)[0][:, :, 0]
    f0 = np.asfortranarray(np.stack([L0 @ z[3 * m + j * n:3 * m + (j + 1) * n] for j in range(m)], axis=1))
    O = orc.ConstantPanel(y, C, bps, tps, theta_os=1.0, theta_ls=ls, KERNEL="Matern")
    assert O.regime == {"CST": 0, "RDM": 1, "GP": 2}[regime]  # CST: one session only (at H > 1 the reference aborts there)
    O.init_constant(theta0, beta0, f0, thr0)
    assert O.draw_fstar(orc.mt()) == 0  # initial draw_fstar on the unseeded generators (src/gpirtMCMC.cpp:227)
    for it in range(iters):
        O.mt_seed_all(it * 10000)
        assert O.sweep(lambda block, h: orc.mt()) == 0
        s = O.state()
        assert np.array_equal(s["theta"], R["theta"][it]), f"sweep {it}"
        assert np.array_equal(s["beta"], R["beta"][..., it])
        assert np.array_equal(s["thr"][:, 1:-1], R["threshold"][:, 1:-1, :, it])
        assert np.array_equal(s["fstar"][:, :, 0], s["fstar"][:, :, -1]) and np.array_equal(s["thr"][:, :, 0], s["thr"][:, :, -1])
        assert abs(O.total_ll() - R["ll"][it]) <= 1e-7 * abs(R["ll"][it])
        assert np.max(np.abs(s["f"] - R["f"][..., it])) < 1e-4 and np.max(np.abs(s["fstar"] - R["fstar"][..., it])) < 2e-3


def test_reference_CST_regime_at_horizon_2_reads_thresholds_out_of_bounds():
    """theta_ls >= 3H with H > 1 (the R default theta_ls = 10 for H <= 3): compute_ll_sparse indexes thresholds(j, c) with j up to m*H - 1 on
    an m-row matrix (src/draw-theta.cpp:139,162 -> :17-19). Armadillo's bounds check (kept in the stand-in) throws inside the
    `omp parallel for` of draw_theta, which terminates the process — so the reference's own code is run in a child process here. This is
    why gpirt_b200 refuses that regime with an error instead of restating it."""
    import subprocess
    import sys

    code = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
from oracle import ref
from tests.problems import make_problem
n, m, C, H = 20, 4, 3, 2
Ps = [make_problem(n, m, C, seed=30 + h, missing=0.0) for h in range(H)]
y = np.stack([P["y"] for P in Ps], axis=2)
z = np.random.default_rng(0).standard_normal(H * (3 * m + n * m))
ref.set_threads(1)
ref.gpirtMCMC(y, np.stack([P["theta0"] for P in Ps], axis=1), 1, 0, 1, np.zeros((3, m)), Ps[0]["beta_prior_sds"], np.zeros((2, n)),
              np.zeros((2, n)), 1.0, 10.0, "Matern", np.stack([Ps[0]["thr0"]] * H, axis=2), 0, z)
print("SURVIVED")
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "SURVIVED" not in r.stdout
    assert "index out of bounds" in r.stderr
