"""CPU tests of the oracle itself: known-answer cases (SURVEY.md 8c — the reference ships none), independent math
(mpmath, numpy, long double), and the reference's RNG discipline."""
import ctypes as C

import mpmath as mp
import numpy as np
import pytest

from oracle import orc
from tests.problems import make_problem, uniform_streams


def test_pnorm_against_mpmath_40_digits():
    mp.mp.dps = 40
    x = np.concatenate([np.linspace(-37, 8.2, 1500), np.random.default_rng(0).normal(size=500) * 2])
    p = orc.pnorm(x)
    worst = max(abs((mp.mpf(float(pi)) - mp.ncdf(mp.mpf(float(xi)))) / mp.ncdf(mp.mpf(float(xi)))) for xi, pi in zip(x, p))
    assert worst < 2e-15
    assert list(orc.pnorm([-np.inf, 0.0, np.inf, -40.0, 9.0])) == [0.0, 0.5, 1.0, 0.0, 1.0]


def test_K_known_answers():
    sds = [3.0, 2.0, 0.5]
    assert orc.K([0.0], [0.0], sds)[0, 0] == 1.0 + 9.0  # theta = 0 -> 1 + s0^2
    x = np.array([-1.5, 0.3, 2.0])
    Kxx = orc.K(x, x, sds)
    assert np.array_equal(Kxx, Kxx.T)
    assert np.allclose(np.diag(Kxx), 1 + 9 + 4 * x ** 2 + 0.25 * x ** 4, rtol=1e-15)  # K(x,x) = 1 + s0^2 + s1^2 x^2 + s2^2 x^4
    want = np.exp(-0.5 * (x[:, None] - x[None, :]) ** 2) + 4 * np.outer(x, x) + 9 + (0.5 * np.outer(x, x)) ** 2
    assert np.allclose(Kxx, want, rtol=1e-15)


def test_K_time_matern_and_rbf():
    ts = np.arange(4.0)
    M = orc.K_time(ts, ts, 1.5, 2.0, [0.0, 0.0], "Matern")
    d = np.abs(ts[:, None] - ts[None, :])
    want = 2.25 * (1 + np.sqrt(5) * d / 2 + 5 * d * d / 12) * np.exp(-np.sqrt(5) * d / 2)
    assert np.allclose(M, want, rtol=1e-14)
    R = orc.K_time(ts, ts, 1.5, 2.0, [0.3, 0.7], "RBF")
    assert np.allclose(R, 2.25 * np.exp(-d * d / 4) + 0.49 * np.outer(ts, ts) + 0.09, rtol=1e-14)
    assert np.all(orc.K_time(ts, ts, 1.0, 1.0, [0, 0], "other") == 0)  # unknown KERNEL: neither branch runs


def test_threshold_transforms_roundtrip_and_edges():
    t = np.array([-np.inf, -0.8, 0.1, 0.9, np.inf])
    d = orc.threshold_to_delta(t)
    assert d[0] == -0.8 and np.allclose(d[1:], np.log([0.9, 0.8]))
    back = orc.delta_to_threshold(d)
    assert np.isneginf(back[0]) and np.isposinf(back[-1]) and np.allclose(back[1:-1], t[1:-1], rtol=1e-15)
    t2 = np.array([-np.inf, 0.25, np.inf])  # C = 2: both loops are empty (src/log-likelihood.cpp:72,83)
    assert list(orc.threshold_to_delta(t2)) == [0.25]
    assert list(orc.delta_to_threshold([0.25])) == [-np.inf, 0.25, np.inf]


def test_ll_bar_sparse_known_answers():
    n = 7
    thr = np.array([-np.inf, 0.0, np.inf])
    y = np.array([1, 2, 1, np.nan, 2, 2, 1.0])
    val = orc.ll_bar_sparse(np.zeros(n), y, np.zeros(n), thr)
    assert abs(val - 6 * np.log(0.5 + 1e-6)) < 1e-15  # g = 0, C = 2: each observed term is log(0.5 + 1e-6)
    assert orc.ll_bar_sparse(np.zeros(n), np.full(n, np.nan), np.zeros(n), thr) == 0.0  # all-missing item
    mp.mp.dps = 40
    rng = np.random.default_rng(3)
    f, mu = rng.standard_normal(50), rng.standard_normal(50)
    thr5 = np.array([-np.inf, -0.84, -0.25, 0.25, 0.84, np.inf])
    y5 = rng.integers(1, 6, 50).astype(float)
    want = mp.mpf(0)
    for i in range(50):
        c = int(y5[i])
        hi = mp.mpf(1) if c == 5 else mp.ncdf(mp.mpf(float(thr5[c])) - (mp.mpf(float(f[i])) + mp.mpf(float(mu[i]))))
        lo = mp.mpf(0) if c == 1 else mp.ncdf(mp.mpf(float(thr5[c - 1])) - (mp.mpf(float(f[i])) + mp.mpf(float(mu[i]))))
        want += mp.log(hi - lo + mp.mpf("1e-6"))
    got = orc.ll_bar_sparse(f, y5, mu, thr5)
    assert abs((mp.mpf(got) - want) / want) < 1e-13


def test_grid_index_bookkeeping():
    assert orc.grid_index(-5.0) == 0 and orc.grid_index(5.0) == 1000
    assert orc.grid_index(-7.3) == 0 and orc.grid_index(11.0) == 1000  # clamp
    assert orc.grid_value(0) == -5.0 and orc.grid_value(1000) == 5.0 and orc.grid_value(500) == 0.0
    for i in range(1001):  # every grid value maps back to its own index
        assert orc.grid_index(orc.grid_value(i)) == i
    assert orc.grid_index(-5.0 + 0.005 + 1e-12) == 1  # just past a half step rounds up
    assert orc.grid_index(-4.9951) == 0


def test_chol_and_double_solve_against_numpy():
    rng = np.random.default_rng(1)
    x = np.sort(rng.uniform(-3, 3, 180))
    A = orc.K(x, x, [0.5, 0.3, 0.1]) + 0.5 * np.eye(180)
    L, info = orc.chol(A)
    assert info == 0 and np.allclose(L, np.linalg.cholesky(A), rtol=1e-11, atol=1e-13) and np.all(np.triu(L, 1) == 0)
    Lb = np.asfortranarray(A.copy())
    assert orc.lib().orc_chol_blocked(Lb.ctypes.data_as(C.c_void_p), C.c_uint64(180)) == 0
    assert np.allclose(Lb, L, rtol=1e-12, atol=1e-14)
    b = rng.standard_normal(180)
    assert np.allclose(orc.double_solve(L, b), np.linalg.solve(A, b), rtol=1e-9)
    bad = np.eye(5)
    bad[3, 3] = -2.0
    assert orc.chol(bad)[1] == 4  # failing leading minor, 1-based


def test_philox_known_answer_and_uniform_range():
    # Random123 known-answer vector for philox4x32-10: counter = key = 0
    assert [hex(v) for v in orc.philox_raw([0, 0, 0, 0], [0, 0])] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(v) for v in orc.philox_raw([0xffffffff] * 4, [0xffffffff] * 2)] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    u = orc.philox_uniform(7, 0, 3, 1, 11, 10001)
    assert u.min() > 0 and u.max() < 1 and abs(u.mean() - 0.5) < 0.02
    z = orc.philox_normal(7, 0, 3, 0, 11, 20000)
    assert abs(z.mean()) < 0.03 and abs(z.std() - 1) < 0.03


def test_ess_deterministic_trace_with_zero_nu():
    """nu = 0: f' = f cos(eps); with u -> 1 only proposals at least as likely as f are accepted, so the bracket
    shrinks towards eps = 0 and the trace is a pure function of the uniform stream (SURVEY 8c iv)."""
    P = make_problem(60, 5, 2, seed=5)
    O = orc.Sampler(P["y"], 2, P["beta_prior_sds"])
    f0 = np.asfortranarray(np.random.default_rng(0).standard_normal((60, 5)))
    O.set_state(theta=P["theta0"], beta=np.zeros((3, 5)), f=f0, thr=P["thr0"])
    O.compute_mu()
    u = np.asfortranarray(np.tile(np.array([[0.999999], [0.5]] + [[0.5]] * 62), (1, 5)))
    O.draw_f(orc.supplied(u=u), nu=np.zeros((60, 5)))
    ev, mg, over = O.diag("f")
    assert not over and np.all(ev >= 2)
    # accepted angle is eps_k of the bisection sequence pi, pi/2 (then towards 0): f' = f cos(eps)
    fnew = O.state()["f"]
    ratio = fnew[0] / f0[0]
    assert np.all(np.abs(ratio) <= 1 + 1e-12)


def test_reference_rng_discipline_single_thread():
    """mt19937 reseeded with iter*10000 per sweep (src/gpirtMCMC.cpp:267); two oracles run the same chain."""
    P = make_problem(40, 6, 3, seed=9)
    outs = []
    for _ in range(2):
        O = orc.Sampler(P["y"], 3, np.full((3, 6), 0.5, order="F"))
        O.set_state(theta=P["theta0"], beta=P["beta0"] * 0.1, f=np.zeros((40, 6)), thr=P["thr0"])
        O.compute_mu()
        assert O.update_cholesky() == 0
        import os
        info, th, ll = O.run_mt(0, 3)
        assert info == 0
        outs.append((th.copy(), ll.copy()))
    if orc.Sampler(P["y"], 3, np.full((3, 6), 0.5, order="F")).num_threads() == 1:
        assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert np.all(np.isfinite(outs[0][1]))
    assert np.all(np.abs(np.round((outs[0][0] + 5) / 0.01) * 0.01 - 5 - outs[0][0]) < 1e-9)  # theta on the grid


def test_full_sweep_supplied_streams_is_reproducible_and_on_grid():
    n, m, C = 50, 8, 4
    P = make_problem(n, m, C, seed=13)
    rng = np.random.default_rng(1)
    srcs = dict(f=orc.supplied(z=rng.standard_normal((n, m)), u=uniform_streams(rng, m)),
                fstar=orc.supplied(z=rng.standard_normal((orc.NGRID, m))),
                theta=orc.supplied(z=rng.standard_normal((1, n)), u=uniform_streams(rng, n)),
                beta=orc.supplied(z=rng.standard_normal((3, m)), u=uniform_streams(rng, m)),
                thr=orc.supplied(z=rng.standard_normal((C - 1, m)), u=uniform_streams(rng, m)))
    states = []
    for _ in range(2):
        O = orc.Sampler(P["y"], C, np.full((3, m), 0.5, order="F"))
        O.set_state(theta=P["theta0"], beta=P["beta0"] * 0.1, f=np.zeros((n, m)), thr=P["thr0"])
        O.compute_mu()
        assert O.update_cholesky() == 0
        assert O.sweep(srcs) == 0
        states.append(O.state())
    for k in states[0]:
        assert np.array_equal(states[0][k], states[1][k])
    th = states[0]["theta"]
    assert all(orc.grid_value(orc.grid_index(t)) == t for t in th)
    thr = states[0]["thr"]
    assert np.all(np.diff(thr[:, 1:C], axis=1) > 0)  # thresholds stay ordered by construction


def test_blas_backend_of_the_timed_baseline_agrees_with_the_oracle_loops():
    """bench.py times the oracle with chol / trsm / syrk / gemv / trsv routed through the scipy wheel's OpenBLAS (oracle/orc_blas.h), as a
    BLAS-linked R would run Armadillo's calls. Same algorithm, different summation order: L agrees norm-wise to kappa * eps, a draw_f with the same
    mt19937 streams makes the same number of proposals on a small well-separated problem, f* agrees to the kappa-limited tolerance."""
    from oracle import orc
    from tests.problems import make_problem

    if orc.find_openblas() is None:
        pytest.skip("no scipy OpenBLAS on this machine")
    n, m, C = 130, 9, 4
    P = make_problem(n, m, C, seed=5)
    sds = np.full((3, m), 0.4, order="F")  # well conditioned: decisions do not hinge on the last bits of L
    out = []
    try:
        for blas in (False, True):
            assert (orc.use_blas(blas, 2) is not None) == blas
            O = orc.Sampler(P["y"], C, sds, P["theta_prior_sds"])
            rng = np.random.default_rng(3)
            O.set_state(theta=np.sort(rng.uniform(-2.5, 2.5, n)), beta=P["beta0"], thr=P["thr0"], f=np.asfortranarray(rng.standard_normal((n, m))))
            O.compute_mu()
            assert O.update_cholesky() == 0
            L = O.L().copy()
            z = np.asfortranarray(rng.standard_normal((n, m)))
            u = np.asfortranarray(rng.random((64, m)))
            O.draw_f(orc.supplied(z=z, u=u))
            zs = np.asfortranarray(rng.standard_normal((orc.NGRID, m)))
            assert O.draw_fstar(orc.supplied(z=zs)) == 0
            out.append((L, O.diag("f")[0].copy(), O.state()["f"].copy(), O.state()["fstar"].copy()))
    finally:
        orc.use_blas(False)
    (L0, e0, f0, s0), (L1, e1, f1, s1) = out
    assert np.linalg.norm(L0 - L1) / np.linalg.norm(L0) < 1e-9
    assert np.array_equal(e0, e1)
    assert np.max(np.abs(f0 - f1)) < 1e-8
    assert np.max(np.abs(s0 - s1)) / np.max(np.abs(s0)) < 1e-6
