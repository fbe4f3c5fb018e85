"""GPU parity: every subsystem of the sweep, called through the C-ABI, against the CPU oracle on the same
seeded inputs with the same supplied nu / z / uniform streams. Tolerances are written at each check."""
import numpy as np
import pytest

import gpirt_b200
from oracle import orc
from tests.problems import make_problem, uniform_streams

pytestmark = pytest.mark.gpu


def _pair(P, seed=1, **kw):
    S = gpirt_b200.Sampler(P["n"], P["m"], P["C"], seed=seed, **kw)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    O = orc.Sampler(P["y"], P["C"], P["beta_prior_sds"], P["theta_prior_sds"])
    return S, O


def test_device_phi_and_log_against_mpmath():
    import mpmath as mp
    from gpirt_b200.api import debug_probit
    mp.mp.dps = 40
    rng = np.random.default_rng(0)
    x = np.concatenate([np.linspace(-9.5, 9.5, 1501), rng.normal(size=1500) * 2.5, [0.0, -8.6, 8.6, -8.6000001, 8.6000001, np.inf, -np.inf]])
    ph, _ = debug_probit(x)
    assert ph[-2] == 1.0 and ph[-1] == 0.0  # Phi(+inf), Phi(-inf): exact, as pnorm
    err = max(abs(mp.mpf(float(p)) - mp.ncdf(mp.mpf(float(xi)))) for xi, p in zip(x[:-2], ph[:-2]))
    assert err < 5e-16  # absolute; R's pnorm (Cody) is ~1e-16. The 1e-6 floor inside the log makes absolute error the relevant one
    z = np.concatenate([10 ** rng.uniform(-6, 0, 2000), 1 + rng.uniform(-1e-6, 1e-6, 200), [1e-6, 1.0, 1.000001, 0.5]])
    _, lg = debug_probit(z)
    assert max(abs(mp.mpf(float(l)) - mp.log(mp.mpf(float(zi)))) for zi, l in zip(z, lg)) < 4e-15


def test_table_phi_of_the_likelihood_pass_against_mpmath():
    """phi_tab (csrc/probit.cuh): Q and the density at 1103 grid points + a 5-term Taylor step with Hermite coefficients, used by k_item_ll and
    k_theta_table. Absolute error against mpmath at 40 digits, including every half-way point between grid values and the ends of the table."""
    import mpmath as mp
    from gpirt_b200.api import debug_phi_tab
    mp.mp.dps = 40
    rng = np.random.default_rng(1)
    half = (np.arange(0, 1103) + 0.5) / 128.0
    x = np.concatenate([np.linspace(-9.5, 9.5, 2001), rng.normal(size=2000) * 2.5, half, -half * 0.99999, [0.0, -8.6, 8.6, 8.609375, 8.62, -8.62, np.inf, -np.inf]])
    ph = debug_phi_tab(x)
    assert ph[-2] == 1.0 and ph[-1] == 0.0
    err = max(abs(mp.mpf(float(p)) - mp.ncdf(mp.mpf(float(xi)))) for xi, p in zip(x[:-2], ph[:-2]))
    assert err < 2.5e-16  # measured 1.2e-16; R's pnorm (Cody) is ~1e-16
    assert np.isnan(debug_phi_tab(np.array([np.nan]))[0])


def test_K_matches_reference_formula():
    rng = np.random.default_rng(0)
    x1, x2 = rng.standard_normal(257) * 2, rng.standard_normal(130) * 2
    sds = np.array([3.0, 2.0, 0.5])
    got, want = gpirt_b200.K(x1, x2, sds), orc.K(x1, x2, sds)
    assert np.max(np.abs(got - want) / np.abs(want)) < 1e-14  # north_star: 1e-10; only exp() can differ (<= 2 ulp)
    assert gpirt_b200.K([0.0], [0.0], sds)[0, 0] == 1.0 + 9.0


@pytest.mark.parametrize("kernel", ["Matern", "RBF"])
def test_K_time(kernel):
    ts = np.arange(6.0)
    got = gpirt_b200.K_time(ts, ts, 1.3, 2.5, [0.2, 0.7], kernel)
    want = orc.K_time(ts, ts, 1.3, 2.5, [0.2, 0.7], kernel)
    assert np.max(np.abs(got - want) / np.abs(want)) < 1e-14


@pytest.mark.parametrize("n", [5, 64, 100, 333, 1001])
def test_chol_well_conditioned_elementwise(n):
    rng = np.random.default_rng(n)
    x = np.sort(rng.uniform(-3, 3, n))
    A = orc.K(x, x, [0.5, 0.3, 0.0]) + np.eye(n) * 0.5  # distinct theta, large jitter: kappa ~ 1e3
    L = gpirt_b200.chol(A)
    Lo, info = orc.chol(A)
    assert info == 0
    assert np.allclose(np.triu(L, 1), 0.0)
    assert np.max(np.abs(L - Lo)) / np.max(np.abs(Lo)) < 1e-10  # north_star tolerance on a well-conditioned fixture
    assert np.max(np.abs(L @ L.T - A)) / np.max(np.abs(A)) < 1e-14


def test_chol_not_positive_definite_is_an_error():
    A = np.eye(70)
    A[65, 65] = -1.0
    with pytest.raises(gpirt_b200.GpirtError) as e:
        gpirt_b200.chol(A)
    assert e.value.code == 4 and "66" in str(e.value)


@pytest.mark.parametrize("n,on_grid", [(100, False), (700, True)])
def test_cholesky_cache_default_priors(n, on_grid):
    """K with the reference's default sds=3 is ill-conditioned (kappa ~ 1e10, duplicated theta when on the grid):
    criterion = backward error <= 1e-14 and norm-wise forward error vs a long-double factor <= 1e-8."""
    P = make_problem(n, 8, 5, seed=3, on_grid=on_grid)
    S, O = _pair(P)
    S.set_state(theta=P["theta0"])
    S.update_cholesky_cache()
    L = S.get_L()
    Kmat = orc.K(P["theta0"], P["theta0"], [3.0, 3.0, 3.0]) + 1e-6 * np.eye(n)
    assert np.max(np.abs(L @ L.T - Kmat)) / np.max(np.abs(Kmat)) < 1e-14
    Lld, info = orc.chol(Kmat, long_double=True)
    assert info == 0
    assert np.linalg.norm(L - Lld) / np.linalg.norm(Lld) < 1e-8


def test_double_solve():
    rng = np.random.default_rng(5)
    n = 300
    x = np.sort(rng.uniform(-3, 3, n))
    A = orc.K(x, x, [0.5, 0.3, 0.0]) + np.eye(n) * 0.5
    L, _ = orc.chol(A)
    X = rng.standard_normal((n, 7))
    got = gpirt_b200.double_solve(L, X)
    want = np.stack([orc.double_solve(L, X[:, k]) for k in range(7)], axis=1)
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-10


@pytest.mark.parametrize("n,m,C", [(100, 37, 2), (257, 50, 5), (64, 9, 7)])
def test_ll_bar_matches_oracle(n, m, C):
    P = make_problem(n, m, C, seed=11)
    S, O = _pair(P)
    f0 = np.asfortranarray(P["rng"].standard_normal((n, m)))
    for X in (S, O):
        X.set_state(theta=P["theta0"], beta=P["beta0"], f=f0, **({"thresholds": P["thr0"]} if X is S else {"thr": P["thr0"]}))
    O.compute_mu()
    per, tot = S.ll_bar()
    want = O.item_ll()
    assert np.max(np.abs(per - want) / np.abs(want)) < 1e-10  # north_star tolerance
    assert abs(tot - O.total_ll()) / abs(tot) < 1e-12


def _set_both(S, O, theta, beta, f, thr, fstar=None):
    S.set_state(theta=theta, beta=beta, f=f, thresholds=thr, fstar=fstar)
    O.set_state(theta=theta, beta=beta, f=f, thr=thr, fstar=fstar)
    O.compute_mu()


@pytest.mark.parametrize("n,m,C", [(100, 40, 2), (300, 33, 5)])
def test_draw_f_same_states_given_same_nu_and_uniforms(n, m, C):
    P = make_problem(n, m, C, seed=21)
    rng = P["rng"]
    S, O = _pair(P)
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    _set_both(S, O, P["theta0"], P["beta0"], f0, P["thr0"])
    nu = np.asfortranarray(rng.standard_normal((n, m)) * 2.0)
    u = uniform_streams(rng, m)
    S.draw_f(u=u, nu=nu)
    O.draw_f(orc.supplied(u=u), nu=nu)
    ev_o, mg_o, over = O.diag("f")
    ev_s, mg_s = S.diag("f")
    assert not over
    assert np.array_equal(ev_s, ev_o), "ESS accepted after a different number of proposals"
    assert mg_o.min() > 1e-9  # the decisions were not near-ties
    got, want = S.get_state()["f"], O.state()["f"]
    assert np.max(np.abs(got - want)) < 1e-12  # same accepted angle: only cos/sin rounding differs


def test_draw_f_with_supplied_z_uses_L():
    n, m, C = 200, 16, 5
    P = make_problem(n, m, C, seed=22)
    rng = P["rng"]
    S, O = _pair(P)
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    theta = np.sort(rng.uniform(-2, 2, n))
    sds = np.full((3, m), 0.3, order="F")
    S.set_priors(sds, None)
    O = orc.Sampler(P["y"], C, sds, P["theta_prior_sds"])
    _set_both(S, O, theta, P["beta0"], f0, P["thr0"])
    z = np.asfortranarray(rng.standard_normal((n, m)))
    u = uniform_streams(rng, m)
    S.update_cholesky_cache()
    assert O.update_cholesky() == 0
    S.draw_f(z=z, u=u)
    O.draw_f(orc.supplied(z=z, u=u))
    assert np.array_equal(S.diag("f")[0], O.diag("f")[0])
    got, want = S.get_state()["f"], O.state()["f"]
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-8  # nu = L z inherits the kappa-limited error of L


@pytest.mark.parametrize("n,m,C", [(150, 30, 2), (120, 45, 5), (700, 21, 5), (90, 12, 2)])  # with and without the per-sweep term table
def test_draw_theta_same_states(n, m, C):
    P = make_problem(n, m, C, seed=31, on_grid=True)
    rng = P["rng"]
    S, O = _pair(P)
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    fstar = np.asfortranarray(rng.standard_normal((orc.NGRID, m)))
    _set_both(S, O, P["theta0"], P["beta0"], f0, P["thr0"], fstar)
    z = np.asfortranarray(rng.standard_normal((1, n)))
    u = uniform_streams(rng, n)
    S.draw_theta(z=z, u=u)
    O.draw_theta(orc.supplied(z=z, u=u))
    assert np.array_equal(S.diag("theta")[0], O.diag("theta")[0])
    assert O.diag("theta")[1].min() > 1e-9
    assert np.array_equal(S.get_state()["theta"], O.state()["theta"])  # grid values: bit-exact
    S.gather_f()
    O.gather_f()
    assert np.array_equal(S.get_state()["f"], O.state()["f"])  # pure indexing: bit-exact


@pytest.mark.parametrize("ls,os_", [(10.0, 1.0), (0.05, 1.0), (2.5, 1.3), (1.5, 0.7)])
def test_draw_theta_of_one_session_in_every_regime(ls, os_):
    """horizon 1 takes the branch theta_ls selects (src/draw-theta.cpp:112 CST, :171 RDM, :206 GP with the 1 x 1 L_time): each with its own
    prior scale, as the oracle restates it (pinned to the reference's own runs in
    tests/test_oracle_vs_ref.py::test_gpirtMCMC_one_session_in_every_theta_regime); non-zero theta_prior_sds."""
    n, m, C = 160, 25, 4
    P = make_problem(n, m, C, seed=33, on_grid=True)
    P["theta_prior_sds"] = np.asfortranarray(np.vstack([np.full((1, n), 0.4), np.zeros((1, n))]))
    rng = P["rng"]
    S = gpirt_b200.Sampler(n, m, C, seed=1, theta_ls=ls, theta_os=os_)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    O = orc.Sampler(P["y"], C, P["beta_prior_sds"], P["theta_prior_sds"], theta_regime=1 if 0.1 < ls < 3.0 else 0, theta_os=os_)
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    fstar = np.asfortranarray(rng.standard_normal((orc.NGRID, m)))
    _set_both(S, O, P["theta0"], P["beta0"], f0, P["thr0"], fstar)
    z = np.asfortranarray(rng.standard_normal((1, n)))
    u = uniform_streams(rng, n)
    S.draw_theta(z=z, u=u)
    O.draw_theta(orc.supplied(z=z, u=u))
    assert np.array_equal(S.diag("theta")[0], O.diag("theta")[0])
    assert np.array_equal(S.get_state()["theta"], O.state()["theta"])
    S.close()


@pytest.mark.parametrize("n,m,C", [(150, 30, 2), (120, 45, 5)])
def test_draw_beta_and_threshold_same_states(n, m, C):
    P = make_problem(n, m, C, seed=41, on_grid=True)
    P["y"][:, 3] = np.nan  # an item nobody answered: beta kept, no draws (src/draw-beta.cpp:73,97-99)
    rng = P["rng"]
    S, O = _pair(P)
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    _set_both(S, O, P["theta0"], P["beta0"], f0, P["thr0"])
    z = np.asfortranarray(rng.standard_normal((3, m)))
    u = uniform_streams(rng, m)
    S.draw_beta(z=z, u=u)
    O.draw_beta(orc.supplied(z=z, u=u))
    ev_s, ev_o = S.diag("beta")[0], O.diag("beta")[0]
    assert np.array_equal(ev_s, ev_o) and ev_s[3] == 0
    got, want = S.get_state()["beta"], O.state()["beta"]
    assert np.array_equal(got[:, 3], P["beta0"][:, 3])
    assert np.max(np.abs(got - want)) < 1e-12
    O.compute_mu()
    z = np.asfortranarray(rng.standard_normal((C - 1, m)))
    u = uniform_streams(rng, m)
    S.draw_threshold(z=z, u=u)
    O.draw_threshold(orc.supplied(z=z, u=u))
    assert np.array_equal(S.diag("thr")[0], O.diag("thr")[0])
    got, want = S.get_state()["thresholds"], O.state()["thr"]
    assert np.all(np.isneginf(got[:, 0])) and np.all(np.isposinf(got[:, C]))  # bookkeeping of the open ends: exact
    assert np.max(np.abs(got[:, 1:C] - want[:, 1:C])) < 1e-12


def test_draw_fstar_against_oracle_well_conditioned():
    n, m, C = 150, 12, 5
    P = make_problem(n, m, C, seed=51)
    rng = P["rng"]
    sds = np.full((3, m), 0.3, order="F")
    S = gpirt_b200.Sampler(n, m, C)
    S.set_data(P["y"])
    S.set_priors(sds, P["theta_prior_sds"])
    O = orc.Sampler(P["y"], C, sds, P["theta_prior_sds"])
    theta = np.sort(rng.uniform(-2.5, 2.5, n))
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    _set_both(S, O, theta, P["beta0"], f0, P["thr0"])
    S.update_cholesky_cache()
    assert O.update_cholesky() == 0
    z = np.asfortranarray(rng.standard_normal((orc.NGRID, m)))
    S.draw_fstar(z=z)
    assert O.draw_fstar(orc.supplied(z=z)) == 0
    got, want = S.get_state(fstar=True)["fstar"], O.state()["fstar"]
    # K^-1 f is kappa-limited: two correct fp64 codes agree to ~kappa*eps, not 1e-10
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-6


def test_supplied_stream_overrun_is_reported():
    n, m, C = 64, 8, 2
    P = make_problem(n, m, C, seed=61)
    S, _ = _pair(P)
    S.set_state(theta=P["theta0"], beta=P["beta0"], f=np.zeros((n, m)), thresholds=P["thr0"])
    u = np.full((2, m), 1e-300, order="F")  # slice height log(1e-300): first proposal is always accepted
    S.draw_f(u=u, nu=np.zeros((n, m)))
    u = np.asfortranarray(np.vstack([np.full((1, m), 1.0 - 1e-16), np.full((1, m), 0.5)]))
    with pytest.raises(gpirt_b200.GpirtError) as e:  # u ~ 1: nothing but the current state is acceptable, stream of 2 runs dry
        S.draw_f(u=u, nu=np.ones((n, m)))
    assert e.value.code == 7


@pytest.mark.parametrize("n", [64, 150, 333, 700])
def test_draw_fstar_block_inverse_solver_matches_substitution(n):
    """solver 1 (explicit L^-1 by parallel block inversion, ragged sizes, several merge levels) against solver 0 (blocked
    forward/back substitution) and the oracle, default priors (kappa(K) ~ 1e10): norm-wise per item column."""
    m, C = 10, 4
    P = make_problem(n, m, C, seed=71, on_grid=True)
    rng = P["rng"]
    f0 = np.asfortranarray(rng.standard_normal((n, m)))
    z = np.asfortranarray(rng.standard_normal((orc.NGRID, m)))
    out = {}
    for solver in (0, 1):
        S = gpirt_b200.Sampler(n, m, C)
        S.set_solver(solver)
        S.set_data(P["y"])
        S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        S.set_state(theta=P["theta0"], beta=P["beta0"], f=f0, thresholds=P["thr0"])
        S.draw_fstar(z=z)
        out[solver] = S.get_state(fstar=True)["fstar"]
        S.close()
    O = orc.Sampler(P["y"], C, P["beta_prior_sds"], P["theta_prior_sds"])
    O.set_state(theta=P["theta0"], beta=P["beta0"], f=f0, thr=P["thr0"])
    O.compute_mu()
    assert O.update_cholesky() == 0
    assert O.draw_fstar(orc.supplied(z=z)) == 0
    want = O.state()["fstar"]
    scale = np.max(np.abs(want), axis=0)
    for solver in (0, 1):
        # f0 is white noise, so K^-1 f0 is as large as kappa allows: two correct fp64 codes differ by ~kappa*eps*|f*| here
        assert np.max(np.max(np.abs(out[solver] - want), axis=0) / scale) < 5e-5, f"solver {solver}"
    assert np.max(np.max(np.abs(out[1] - out[0]), axis=0) / scale) < 5e-5


def test_overlapped_sweeps_equal_serial_sweeps_bit_for_bit():
    """gpirt_sweep with the theta-only chain on the side stream must give exactly the state of the serial order (same Philox streams)."""
    n, m, C = 300, 40, 5
    P = make_problem(n, m, C, seed=81)
    states = []
    for overlap in (1, 0):
        S = gpirt_b200.Sampler(n, m, C, seed=99)
        S.set_overlap(overlap)
        S.set_data(P["y"])
        S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
        S.init_draw(draw_beta=False, draw_f=True)
        S.sweep(4)
        S.sync()
        st = S.get_state(fstar=True)
        st["ll"] = S.total_ll()
        st["L"] = S.get_L()
        states.append(st)
        S.close()
    for k in ("theta", "beta", "f", "thresholds", "fstar", "L"):
        assert np.array_equal(states[0][k], states[1][k]), k
    assert states[0]["ll"] == states[1]["ll"]


def test_profiled_sweeps_leave_the_chain_alone_and_the_timeline_is_ordered():
    """gpirt_profile_*: the serial order (level 1) and the overlapped schedule with events around every group (level 2) give the chain of unprofiled
    sweeps; a reset forgets the sections recorded so far (collected or not); the timeline lists each group of each recorded sweep with begin <= end."""
    n, m, C = 300, 40, 5
    P = make_problem(n, m, C, seed=81)
    states = []
    for level in (0, 1, 2):
        S = gpirt_b200.Sampler(n, m, C, seed=99)
        S.set_data(P["y"])
        S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
        S.init_draw(draw_beta=False, draw_f=True)
        S.profile(level)
        S.sweep(2)
        S.profile_reset()
        S.sweep(3)
        if level == 1:
            sec = S.profile_read()
            # every group once per recorded sweep: the two sweeps before the reset are gone, and the shared f* terms are formed in each recorded sweep
            for name in ("gemm_LZ", "ess_f", "gemm_FtW", "ess_theta", "ess_beta", "ess_thr", "item_ll", "chol_K", "trtri", "gemm_tmp", "gemm_W", "chol_Kpost"):
                assert sec[name][1] == 3, (name, sec[name])
                assert sec[name][0] > 0.0
            assert not any(k.startswith("wait_") for k in sec)
        if level == 2:
            tl = S.profile_timeline()
            names = [t[0] for t in tl]
            assert names.count("gemm_LZ") == 3 and names.count("chol_K") == 3 and names.count("chol_Kpost") == 3
            assert all(0.0 <= b <= e for _, b, e in tl)
            lz = [b for name, b, _ in tl if name == "gemm_LZ"]
            assert lz == sorted(lz)
            assert S.profile_timeline() == []  # collected
        S.profile(0)
        S.sync()
        st = S.get_state(fstar=True)
        st["ll"] = S.total_ll()
        states.append(st)
        S.close()
    for st in states[1:]:
        for k in ("theta", "beta", "f", "thresholds", "fstar"):
            assert np.array_equal(states[0][k], st[k]), k
        assert states[0]["ll"] == st["ll"]


def test_fp32_screening_term_stays_inside_its_error_bound():
    """cell_ll32 (csrc/probit.cuh): |t32 - t64| <= e32 over random and adversarial cells (tails, narrow bins, edge categories)."""
    from gpirt_b200.api import debug_cell32

    rng = np.random.default_rng(5)
    n = 1_000_000
    g = rng.normal(0, 3, n)
    lo = rng.normal(0, 2, n)
    hi = lo + np.exp(rng.normal(-0.5, 2.0, n))  # bins from 1e-4 to 50 wide
    kind = rng.integers(0, 4, n)
    lo = np.where(kind == 1, -np.inf, lo)
    hi = np.where(kind == 2, np.inf, hi)
    far = kind == 3  # deep tails on either side
    g = np.where(far, g + rng.choice([-1.0, 1.0], n) * rng.uniform(4, 14, n), g)
    t32, e32, t64 = debug_cell32(g, lo, hi)
    assert np.all(np.isfinite(t64)) and np.all(np.isfinite(e32))
    ratio = np.abs(t32 - t64) / e32
    assert ratio.max() <= 0.6, f"observed error reaches {ratio.max():.2f} of the bound"  # the bound carries a 2x margin by design
    assert np.median(e32) < 5e-5  # and it is tight enough to decide: ~0.03 per 2000 cells against margins of O(1..1000)


@pytest.mark.parametrize("n,m,C", [(300, 40, 5), (257, 33, 2), (500, 20, 7)])
def test_screened_sweeps_equal_unscreened_sweeps_bit_for_bit(n, m, C):
    """fp32 screening on vs off: the same decisions, so identical states after several production sweeps; and the screen decides most
    proposals on its own."""
    P = make_problem(n, m, C, seed=91)
    states, exact, evals = [], None, None
    for screen in (1, 0, 2):  # 2: screening plus early rejection on partial sums (off by default, samplers.cu)
        S = gpirt_b200.Sampler(n, m, C, seed=5)
        S.set_screen(screen)
        S.set_data(P["y"])
        S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
        S.init_draw(draw_beta=False, draw_f=True)
        S.sweep(5)
        S.sync()
        st = S.get_state(fstar=True)
        st["ll"] = S.total_ll()
        st["evals"] = {b: S.diag(b)[0] for b in ("f", "beta", "thr", "theta")}
        if screen == 1:
            exact = {b: S.exact_evals(b) for b in ("f", "beta", "thr", "theta")}
        states.append(st)
        S.close()
    for other in (1, 2):
        for k in ("theta", "beta", "f", "thresholds", "fstar"):
            assert np.array_equal(states[0][k], states[other][k]), (other, k)
        assert states[0]["ll"] == states[other]["ll"]
        for b in ("f", "beta", "thr", "theta"):
            assert np.array_equal(states[0]["evals"][b], states[other]["evals"][b]), (other, b)
    for b in ("f", "beta", "thr"):
        assert np.array_equal(states[0]["evals"][b], states[1]["evals"][b])
        ev = states[0]["evals"][b]
        # proposals that needed the fp64 pass (the current state is screened too: an item's first undecided proposal costs one extra exact evaluation)
        frac = np.maximum(exact[b][ev > 0] - 1, 0).sum() / max(1, (ev[ev > 0] - 1).sum())
        assert frac < 0.25, f"{b}: {frac:.2f} of the proposals fell back to fp64"
    # theta step: the current state is screened too (an undecided proposal costs the current state's exact evaluation once, plus its own)
    ev, ex = states[0]["evals"]["theta"], exact["theta"]
    assert np.array_equal(ev, states[1]["evals"]["theta"])
    frac = np.maximum(ex - 1, 0).sum() / max(1, (ev - 1).sum())
    assert frac < 0.25, f"theta: {frac:.2f} of the proposals fell back to fp64"


def test_full_size_chain_is_deterministic_across_overlap_and_screening():
    """At the C3 size (n=2000, m=5000) the side-stream chain really runs under the item kernels, with delayed CTAs: the chain must
    still be bit-identical to the serial, unscreened order (this caught an in-place race in the Cholesky panel kernel)."""
    n, m, C, sweeps = 2000, 5000, 5, 6
    P = make_problem(n, m, C, seed=20261104 + 3, dtype="i8")
    out = []
    for overlap, screen in ((0, 0), (1, 1), (1, 1)):
        S = gpirt_b200.Sampler(n, m, C, seed=20261104)
        S.set_overlap(overlap)
        S.set_screen(screen)
        S.set_data(P["y"])
        S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
        S.init_draw(draw_beta=False, draw_f=True)
        S.sweep(sweeps)
        S.sync()
        out.append((S.total_ll(), S.get_state(f=False)["theta"], S.get_state(f=False)["beta"]))
        S.close()
    for k in (1, 2):
        assert out[k][0] == out[0][0] and np.array_equal(out[k][1], out[0][1]) and np.array_equal(out[k][2], out[0][2])


@pytest.mark.gpu
@pytest.mark.parametrize("screen", [1, 0])
@pytest.mark.parametrize("n,m,C", [(300, 1500, 5), (150, 2100, 3), (64, 1024, 2), (80, 37, 4)])  # ragged and exact multiples of the 1024 virtual lanes
def test_theta_kernel_gives_the_same_bits_with_256_and_1024_threads_per_respondent(n, m, C, screen):
    """The theta step runs 256 threads per respondent, or 1024 on a rank with few respondents (item shards); both sum an exact evaluation in the
    same order, so a chain must not depend on the choice — with and without the per-sweep term table, with and without fp32 screening (the
    screening sums may differ in the last bits between the two; decisions cannot, and the recorded margin is then a lower bound, compared to 1e-9)."""
    import gpirt_b200 as g
    from gpirt_b200 import _lib

    P = make_problem(n, m, C, seed=13)
    states = {}
    try:
        for threads in (256, 1024):
            _lib.check(_lib.lib().gpirt_debug_theta_threads(threads))
            S = g.Sampler(n, m, C, seed=5)
            S.set_screen(screen)
            S.set_data(P["y"])
            S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
            S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
            S.init_draw(draw_beta=False, draw_f=True)
            S.sweep(3)
            st = S.get_state(fstar=True)
            st["ll"], st["ev"], st["mg"] = S.total_ll(), S.diag("theta")[0], S.diag("theta")[1]
            states[threads] = st
    finally:
        _lib.lib().gpirt_debug_theta_threads(0)
    a, b = states[256], states[1024]
    for k in ("theta", "beta", "thresholds", "f", "fstar", "ev"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["mg"], b["mg"]) if not screen else np.allclose(a["mg"], b["mg"], rtol=1e-9, atol=1e-9)
    assert a["ll"] == b["ll"]


@pytest.mark.gpu
def test_chol_large_matrix_two_level_blocking_backward_error():
    """n >= 4096 switches the blocked Cholesky to two levels (K = 256 trailing updates, config C4 has n = 10 000): backward error and agreement
    with LAPACK on a well-conditioned covariance."""
    import gpirt_b200 as g

    n = 4200  # not a multiple of 64 or 256
    rng = np.random.default_rng(4)
    th = np.clip(rng.standard_normal(n) * 1.5, -5, 5)
    A = g.K(th, th, np.array([0.8, 0.5, 0.1])) + 0.5 * np.eye(n)
    L = g.chol(A)
    assert np.all(np.triu(L, 1) == 0)
    R = L @ L.T - A
    assert np.max(np.abs(R)) <= 2e-14 * np.max(np.abs(A)) * 8
    Lref = np.linalg.cholesky(A)
    assert np.max(np.abs(L - Lref)) < 1e-11


@pytest.mark.gpu
def test_cholesky_cache_with_look_ahead_equals_the_plain_two_level_factorisation_bit_for_bit():
    """A sampler with n >= 4096 factorises K with the look-ahead schedule (the bulk of every K = 256 trailing update on a helper stream beside the
    next panels, linalg.cuh CholAux); gpirt_chol() on the same matrix runs the plain two-level order. Same tiles, same k ranges: identical L."""
    import gpirt_b200 as g

    n, m, C = 4500, 8, 3
    P = make_problem(n, m, C, seed=41, on_grid=True)
    S = g.Sampler(n, m, C)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.update_cholesky_cache()
    L = S.get_L()
    th = S.get_state(f=False)["theta"]
    A = g.K(th, th, P["beta_prior_sds"][:, 0]) + 1e-6 * np.eye(n)
    assert np.array_equal(L, g.chol(A))
    assert np.max(np.abs(L @ L.T - A)) <= 1e-13 * np.max(np.abs(A)) * 8


@pytest.mark.gpu
@pytest.mark.parametrize("n,m,C,name", [(100, 418, 2, "C1 senate116 shape"), (1500, 16, 5, "C2 SDO shape"), (333, 75, 7, "ragged"),
                                        (90, 14, 3, "holes"), (220, 18, 21, "GPIRT_MAX_C categories"),
                                        # the kernel variants BASELINE configs 3 / 4 / 5 select (samplers.cu item_warps, linalg.cu two-level blocking):
                                        (2000, 96, 5, "C3 respondent count: 8-warp item kernels, 32 Cholesky panels"),
                                        (5000, 24, 2, "C5 respondent count: 16-warp item kernels, two-level Cholesky with look-ahead inside a sweep"),
                                        (4500, 16, 7, "C4 category count at n > 4096")])
def test_sweeps_at_baseline_config_shapes_match_oracle_on_the_same_philox_streams(n, m, C, name):
    """Whole sweeps (gpirt_sweep: overlapped schedule, fp32 screening, block inverse) against the oracle's sweep on the SAME Philox streams, at the
    shapes of BASELINE.json's configs 1 and 2 (synthetic responses; the .rda front-end is out of scope) with the reference's default priors
    (beta_prior_sds = 3: kappa(K) ~ 1e10, so f / f* are compared norm-wise per item and theta exactly)."""
    import gpirt_b200 as g

    P = make_problem(n, m, C, seed=77, missing=0.06)
    if name == "holes":  # a respondent without answers, an item nobody answered, an item with one answer (the reference's own run on such
        P["y"][5, :] = np.nan  # an input: tests/test_oracle_vs_ref.py::test_gpirtMCMC_with_a_respondent_and_an_item_without_answers)
        P["y"][:, 2] = np.nan
        P["y"][:, 4] = np.nan
        P["y"][11, 4] = 2.0
    seed, chain, sweeps = 2024, 1, 2
    S = g.Sampler(n, m, C, seed=seed, chain=chain)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    O = orc.Sampler(P["y"], C, P["beta_prior_sds"], P["theta_prior_sds"])
    rng = np.random.default_rng(1)
    O.set_state(theta=P["theta0"], beta=P["beta0"], thr=P["thr0"])
    O.compute_mu()  # the oracle's sweep, like the reference's loop body, starts from mu and L of the current state (src/gpirtMCMC.cpp:136,308-314)
    assert O.update_cholesky() == 0
    f0 = np.asfortranarray(O.L() @ rng.standard_normal((n, m)))  # f0_j = L z_j as the reference initialises it (:158-162); white noise is not a GP draw
    S.set_state(theta=P["theta0"], beta=P["beta0"], f=f0, thresholds=P["thr0"])
    O.set_state(f=f0)
    for it in range(sweeps):
        S.set_iter(it)
        S.sweep(1)
        assert O.sweep(orc.philox(seed, chain, it)) == 0
        a, b = S.get_state(fstar=True), O.state()
        for blk in ("f", "theta", "beta", "thr"):
            assert np.array_equal(S.diag(blk)[0], O.diag(blk)[0]), f"sweep {it}: block {blk} made a different number of proposals"
        assert np.array_equal(a["theta"], b["theta"])
        assert np.max(np.abs(a["beta"] - b["beta"])) < 1e-9 and np.max(np.abs(a["thresholds"][:, 1:-1] - b["thr"][:, 1:-1])) < 1e-9
        colrel = lambda x, y: float(np.max(np.max(np.abs(x - y), axis=0) / np.maximum(1.0, np.max(np.abs(y), axis=0))))
        # f* also extrapolates to |theta*| = 5 from respondents inside about +-2.7 (values ~100): K^-1 f is kappa-limited there
        # (kappa grows with n: at n = 2000 and n = 4500 two correct fp64 codes were measured 3.1e-5 apart on f*, f and every decision still agreeing)
        assert colrel(a["f"], b["f"]) < 4e-6 and colrel(a["fstar"], b["fstar"]) < (3e-5 if n < 2000 else 1e-4)
        assert abs(S.total_ll() - O.total_ll()) <= 1e-7 * abs(O.total_ll())


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [256, 1024])
def test_draw_theta_over_a_wide_item_row_matches_oracle(threads):
    """Config 4's theta step: m = 50 000 items per respondent row. With m > 24 576 the 256-thread kernel no longer stages the respondent's row of y
    in shared memory (samplers.cu launch_ess_theta, y_in_smem == false) and with m > 12 000 the per-sweep term table is the unscreened fp64 one
    (capi.cu do_draw_theta): both branches against the oracle on supplied streams, theta bit-exact, identical proposal counts
    (src/draw-theta.cpp:4-84,112-170)."""
    from gpirt_b200 import _lib

    n, m, C = 48, 41000, 2  # 1024 threads: 41 008 bytes of dynamic + ~9 KB of static shared memory, i.e. above the 48 KB default (8 shards of C3 land here)
    P = make_problem(n, m, C, seed=35, on_grid=True)
    rng = P["rng"]
    try:
        _lib.check(_lib.lib().gpirt_debug_theta_threads(threads))
        S, O = _pair(P)
        f0 = np.zeros((n, m), order="F")
        fstar = np.asfortranarray(rng.standard_normal((orc.NGRID, m)) * 0.3)
        _set_both(S, O, P["theta0"], P["beta0"] * 0.1, f0, P["thr0"], fstar)
        z = np.asfortranarray(rng.standard_normal((1, n)))
        u = uniform_streams(rng, n, stride=96)
        S.draw_theta(z=z, u=u)
        O.draw_theta(orc.supplied(z=z, u=u))
        assert np.array_equal(S.diag("theta")[0], O.diag("theta")[0])
        assert O.diag("theta")[1].min() > 1e-9
        assert np.array_equal(S.get_state(f=False)["theta"], O.state()["theta"])
        S.close()
    finally:
        _lib.lib().gpirt_debug_theta_threads(0)
