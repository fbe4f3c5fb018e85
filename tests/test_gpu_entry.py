"""The two reference entry points through the C-ABI (gpirt_mcmc <- gpirtMCMC(), src/gpirtMCMC.cpp:25-406; gpirt_recover_fstar <-
recover_fstar(), src/recover_fstar.cpp:8-93) against the same chain driven step by step through the subsystem calls, and the
device-side draw store behind them. Everything here is bit-for-bit: the entry points add bookkeeping (burn-in, thinning, sample
index, layouts of the kept draws), no arithmetic."""
import numpy as np
import pytest

from tests.problems import make_problem

pytestmark = pytest.mark.gpu


def stepwise_chain(P, seed, total, keep):
    """The chain gpirt_mcmc runs (gpirt_b200/csrc/entry.cu), one subsystem call at a time; returns the states after the sweeps in `keep`."""
    import gpirt_b200 as g

    n, m, C = P["n"], P["m"], P["C"]
    S = g.Sampler(n, m, C, seed=seed)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.set_iter(0xFFFFFFFF)
    S.init_draw(draw_beta=False, draw_f=True)
    S.set_iter(0)
    out = {}
    for it in range(total):
        S.draw_f()
        S.draw_fstar()
        S.draw_theta()
        S.gather_f()
        S.draw_beta()
        S.update_cholesky_cache()
        S.draw_threshold()
        _, ll = S.ll_bar()
        S.set_iter(it + 1)
        if it in keep:
            st = S.get_state(fstar=True)
            st["ll"] = ll
            out[it] = st
    return out


@pytest.mark.parametrize("n,m,C,sample,burn,thin", [(120, 17, 4, 7, 1, 2), (90, 11, 2, 5, 0, 1), (64, 9, 5, 6, 3, 3)])
def test_gpirtMCMC_equals_the_stepwise_chain(n, m, C, sample, burn, thin):
    import gpirt_b200 as g

    P = make_problem(n, m, C, seed=5 + n)
    seed = 77
    res = g.gpirtMCMC(P["y"], sample, burn, THIN=thin, CHAIN=1, beta_prior_sds=P["beta_prior_sds"], theta_prior_sds=P["theta_prior_sds"],
                      theta_init=P["theta0"], thresholds=P["thr0"], SEED=seed, beta_init=P["beta0"], store_f=True, store_fstar=True)[0]
    total = sample + burn
    n_samples = int(sample / thin)  # src/gpirtMCMC.cpp:48
    kept = [it for it in range(total) if it >= burn and it % thin == 0 and (it - burn) // thin < n_samples]  # :334-335
    assert res["theta"].shape == (n_samples, n, 1) and len(res["beta"]) == n_samples
    ref = stepwise_chain(P, seed, total, set(kept))
    stored = sorted({(it - burn) // thin for it in kept})
    for it in kept:
        s = (it - burn) // thin  # a later sweep with the same index overwrites an earlier one, as in the reference
        if it != max(k for k in kept if (k - burn) // thin == s):
            continue
        st = ref[it]
        assert np.array_equal(res["theta"][s, :, 0], st["theta"])
        assert np.array_equal(res["beta"][s][:, :, 0], st["beta"])
        assert np.array_equal(res["threshold"][s][:, :, 0], st["thresholds"])
        assert np.array_equal(res["f"][s][:, :, 0], st["f"])
        assert np.array_equal(res["fstar"][s][:, :, 0], st["fstar"])
        assert res["ll"][s] == st["ll"]
    assert stored == list(range(len(stored)))


def test_draw_store_in_several_flushes_equals_get_state():
    import gpirt_b200 as g

    n, m, C = 150, 23, 5
    P = make_problem(n, m, C, seed=31)
    S = g.Sampler(n, m, C, seed=9)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True)
    n_samples = 5
    out = dict(theta=np.zeros((n_samples, n, 1), order="F"), beta=np.zeros((3, m, 1, n_samples), order="F"),
               threshold=np.zeros((m, C + 1, 1, n_samples), order="F"), ll=np.zeros(n_samples), f=np.zeros((n, m, 1, n_samples), order="F"),
               fstar=np.zeros((g.NGRID, m, 1, n_samples), order="F"))
    assert S.store_sample_bytes(True, True) == 8 * (n + 3 * m + (C + 1) * m + 1 + n * m + g.NGRID * m)
    S.store_open(2, store_f=True, store_fstar=True)
    states, first = [], 0
    for s in range(n_samples):
        S.sweep(1)
        S.store_push()
        st = S.get_state(fstar=True)  # synchronises; the store must not depend on that
        st["ll"] = S.total_ll()
        states.append(st)
        cnt, cap = S.store_count()
        if cnt == cap or s == n_samples - 1:
            S.store_flush(out, first)
            first += cnt
            assert S.store_count()[0] == 0
    with pytest.raises(g.GpirtError):
        for _ in range(3):
            S.store_push()  # capacity 2
    S.store_close()
    for s, st in enumerate(states):
        assert np.array_equal(out["theta"][s, :, 0], st["theta"])
        assert np.array_equal(out["beta"][:, :, 0, s], st["beta"])
        assert np.array_equal(out["threshold"][:, :, 0, s], st["thresholds"])
        assert np.array_equal(out["f"][:, :, 0, s], st["f"])
        assert np.array_equal(out["fstar"][:, :, 0, s], st["fstar"])
        assert out["ll"][s] == st["ll"]


def test_recover_fstar_equals_draw_f_then_draw_fstar():
    import gpirt_b200 as g

    n, m, C = 110, 13, 4
    P = make_problem(n, m, C, seed=3)
    rng = np.random.default_rng(0)
    f = np.asfortranarray(rng.standard_normal((n, m)))
    beta2 = np.asfortranarray(rng.standard_normal((2, m)))
    bps2 = np.full((2, m), 0.5, order="F")
    seed = 12345
    got = g.recover_fstar(seed, f, P["y"], P["theta0"], beta2, P["thr0"], beta_prior_sds=bps2)["fstar"][:, :, 0]
    S = g.Sampler(n, m, C, seed=seed, p_beta=2)
    S.set_data(P["y"])
    sds3 = np.zeros((3, m), order="F")
    sds3[:2] = bps2  # K() reads a third row the 2 x m matrix does not have (src/covariance-function.cpp:12-14); taken as 0
    S.set_priors(sds3, None)
    S.set_state(theta=P["theta0"], beta=beta2, f=f, thresholds=P["thr0"])
    S.set_iter(0)
    S.draw_f()
    S.draw_fstar()
    assert np.array_equal(got, S.get_state(fstar=True)["fstar"])


def _recover_inputs(n, m, C, H, seed):
    rng = np.random.default_rng(seed)
    Ps = [make_problem(n, m, C, seed=seed + h) for h in range(H)]
    y = np.asfortranarray(np.stack([p["y"] for p in Ps], axis=2))
    theta = np.asfortranarray(np.stack([p["theta0"] for p in Ps], axis=1))
    thr = np.asfortranarray(np.stack([p["thr0"] for p in Ps], axis=2))
    f = np.asfortranarray(rng.standard_normal((n, m, H)))
    beta = np.asfortranarray(rng.standard_normal((2, m, H)))
    sds3 = np.zeros((3, m), order="F")
    sds3[:2] = 0.5
    return y, theta, thr, f, beta, sds3


def test_recover_fstar_over_a_panel_runs_every_session():
    """horizon > 1, constant_IRF = 0: draw_f / draw_fstar session by session (src/draw-f.cpp:141-179, src/draw-fstar.cpp:127-168)."""
    import gpirt_b200 as g

    n, m, C, H, seed = 70, 9, 3, 3, 77
    y, theta, thr, f, beta, sds3 = _recover_inputs(n, m, C, H, 11)
    got = g.recover_fstar(seed, f, y, theta, beta, thr, beta_prior_sds=sds3[:2])["fstar"]
    assert got.shape == (g.NGRID, m, H)
    for h in range(H):
        S = g.Sampler(n, m, C, seed=seed, chain=h, p_beta=2)
        S.set_data(y[:, :, h])
        S.set_priors(sds3, None)
        S.set_state(theta=theta[:, h], beta=beta[:, :, h], f=f[:, :, h], thresholds=thr[:, :, h])
        S.set_iter(0)
        S.draw_f()
        S.draw_fstar()
        assert np.array_equal(got[:, :, h], S.get_state(fstar=True)["fstar"])
        S.close()
    assert not np.array_equal(got[:, :, 0], got[:, :, 1])


def test_recover_fstar_with_constant_IRF_shares_one_curve():
    """constant_IRF = 1: stacked draw_f (src/draw-f.cpp:84-139), f interpolated to 100 inducing points, one f* for all sessions (src/draw-fstar.cpp:58-125)."""
    import gpirt_b200 as g

    n, m, C, H, seed = 60, 7, 3, 2, 5
    y, theta, thr, f, beta, sds3 = _recover_inputs(n, m, C, H, 23)
    got = g.recover_fstar(seed, f, y, theta, beta, thr, beta_prior_sds=sds3[:2], constant_IRF=1)["fstar"]
    for h in range(1, H):
        assert np.array_equal(got[:, :, h], got[:, :, 0])
    assert np.isfinite(got).all()
    ths = np.concatenate([theta[:, h] for h in range(H)])
    ys = np.asfortranarray(np.concatenate([y[:, :, h] for h in range(H)], axis=0))
    fs = np.asfortranarray(np.concatenate([f[:, :, h] for h in range(H)], axis=0))
    mu = np.asfortranarray(np.concatenate([beta[0, :, h][None, :] + theta[:, h][:, None] * beta[1, :, h][None, :] for h in range(H)], axis=0))
    T = g.Sampler(n * H, m, C, seed=seed, p_beta=2)
    T.set_data(ys)
    T.set_priors(sds3, None)
    T.set_state(theta=ths, beta=beta[:, :, 0], f=fs, thresholds=thr[:, :, 0])
    T.set_mean_rows(mu)
    T.set_iter(0)
    T.draw_f()
    fn = T.get_state()["f"]
    lo, hi = float(ths.min()), float(ths.max())
    tind = lo + np.arange(100, dtype=np.float64) * ((hi - lo) / 99)
    tind[-1] = hi
    find = g.interp1(ths, fn, tind)
    I = g.Sampler(100, m, C, seed=seed, p_beta=2)
    I.set_data(np.full((100, m), np.nan, order="F"))
    I.set_priors(sds3, None)
    I.set_state(theta=tind, beta=beta[:, :, 0], f=find)
    I.set_iter(0)
    I.draw_fstar()
    assert np.array_equal(got[:, :, 0], I.get_state(fstar=True)["fstar"])
    T.close()
    I.close()
    # one session: the reference runs the same branch (the f* draw then goes through the inducing points instead of the n respondents)
    one = g.recover_fstar(seed, f[:, :, :1], y[:, :, :1], theta[:, :1], beta[:, :, :1], thr[:, :, :1], beta_prior_sds=sds3[:2], constant_IRF=1)["fstar"]
    plain = g.recover_fstar(seed, f[:, :, :1], y[:, :, :1], theta[:, :1], beta[:, :, :1], thr[:, :, :1], beta_prior_sds=sds3[:2])["fstar"]
    assert one.shape == plain.shape == (g.NGRID, m, 1) and np.isfinite(one).all() and not np.array_equal(one, plain)


# ---- the R-package shim (integration/r-package/src) driven without R, tests/rshim ------------------------------------
def _rshim():
    import ctypes as C
    import os

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "rshim", "_build", "librshim_test.so")
    if not os.path.exists(path):
        pytest.skip("tests/rshim/_build/librshim_test.so not built (make -C tests/rshim, or __graft_entry__.build())")
    L = C.CDLL(path)
    L.shim_last_error.restype = C.c_char_p
    L.shim_normals_used.restype = C.c_uint64
    return L


def test_r_shim_gpirtMCMC_returns_what_the_c_abi_returns():
    """gpirtMCMC() of the R shim (stand-in Rcpp/Armadillo types, R's RNG = a supplied stream) against gpirt_b200.gpirtMCMC with the initial
    values the shim must have drawn: beta0 = mean + sd z in (item, coefficient) order, then n normals per item for f0 = L z."""
    import ctypes as C

    import gpirt_b200 as g

    L = _rshim()
    n, m, Ccat, sample, burn, thin = 80, 12, 4, 6, 2, 2
    P = make_problem(n, m, Ccat, seed=41)
    rng = np.random.default_rng(8)
    z = rng.standard_normal(3 * m + n * m + 5)
    unif = 0.3712
    bpm = np.asfortranarray(rng.standard_normal((3, m)) * 0.1)
    bps, tps = P["beta_prior_sds"], P["theta_prior_sds"]
    tpm = np.zeros((2, n), order="F")
    ns = sample // thin
    out = dict(theta=np.zeros((ns, n, 1), order="F"), beta=np.zeros((3, m, 1, ns), order="F"), thr=np.zeros((m, Ccat + 1, 1, ns), order="F"),
               ll=np.zeros(ns), f=np.zeros((n, m, 1, ns), order="F"), fstar=np.zeros((g.NGRID, m, 1, ns), order="F"))
    y = np.asfortranarray(P["y"])
    th0 = np.ascontiguousarray(P["theta0"] * 1.6)  # beyond +-5 somewhere: the clamp of src/gpirtMCMC.cpp:133 happens behind the ABI
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    u64 = C.c_uint64
    rc = L.shim_gpirtMCMC(u64(n), u64(m), u64(1), u64(Ccat), p(y), p(th0), sample, burn, thin, p(bpm), p(bps), p(tpm), p(tps), C.c_double(1.0),
                          C.c_double(10.0), b"Matern", p(P["thr0"]), 0, p(z), u64(z.size), C.c_double(unif), p(out["theta"]), p(out["beta"]),
                          p(out["thr"]), p(out["ll"]), p(out["f"]), p(out["fstar"]))
    assert rc == 0, L.shim_last_error().decode()
    assert L.shim_normals_used() == 3 * m + n * m  # src/gpirtMCMC.cpp:149-163: 3 per item, then n per item
    beta0 = np.asfortranarray(bpm + bps * z[:3 * m].reshape(m, 3).T)
    z0 = np.asfortranarray(z[3 * m:3 * m + n * m].reshape(m, n).T)
    ref = g.gpirtMCMC(y, sample, burn, THIN=thin, CHAIN=1, beta_prior_means=bpm, beta_prior_sds=bps, theta_prior_sds=tps, theta_init=th0,
                      thresholds=P["thr0"], SEED=int(unif * 4294967296.0), beta_init=beta0, f_init_z=z0, store_f=True, store_fstar=True)[0]
    assert np.array_equal(out["theta"], ref["theta"]) and np.array_equal(out["ll"], ref["ll"])
    for s in range(ns):
        assert np.array_equal(out["beta"][..., s], ref["beta"][s]) and np.array_equal(out["thr"][..., s], ref["threshold"][s])
        assert np.array_equal(out["f"][..., s], ref["f"][s]) and np.array_equal(out["fstar"][..., s], ref["fstar"][s])
    # and f_init_z is what it says: the first sweep starts from f0 = L z
    S = g.Sampler(n, m, Ccat, seed=1)
    S.set_priors(bps, tps)
    S.set_state(theta=th0)
    Lk = g.chol(g.K(np.clip(th0, -5, 5), np.clip(th0, -5, 5), bps[:, 0]) + 1e-6 * np.eye(n))
    S.init_f_from_z(z0)
    f0 = S.get_state(theta=False, beta=False, thresholds=False)["f"]
    assert np.max(np.abs(f0 - Lk @ z0)) <= 1e-9 * np.max(np.abs(f0))


def test_r_shim_recover_fstar_and_error_path():
    import ctypes as C

    import gpirt_b200 as g

    L = _rshim()
    n, m, Ccat = 70, 9, 3
    P = make_problem(n, m, Ccat, seed=2)
    rng = np.random.default_rng(1)
    f = np.asfortranarray(rng.standard_normal((n, m)))
    beta2 = np.asfortranarray(rng.standard_normal((2, m)))
    bpm, bps = np.zeros((2, m), order="F"), np.full((2, m), 0.5, order="F")
    y = np.asfortranarray(P["y"])
    th = np.ascontiguousarray(P["theta0"])
    out = np.zeros((g.NGRID, m), order="F")
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    u64 = C.c_uint64
    rc = L.shim_recover_fstar(321, u64(n), u64(m), u64(1), u64(Ccat), p(f), p(y), p(th), p(beta2), p(P["thr0"]), p(bpm), p(bps), 0, p(out))
    assert rc == 0, L.shim_last_error().decode()
    assert np.array_equal(out, g.recover_fstar(321, f, y, th, beta2, P["thr0"], beta_prior_sds=bps)["fstar"][:, :, 0])
    # a library error becomes Rcpp::stop(message) -> an R error; here a C++ exception carrying gpirt_last_error()
    thr_bad = np.zeros((m, 31), order="F")
    rc = L.shim_recover_fstar(321, u64(n), u64(m), u64(1), u64(30), p(f), p(y), p(th), p(beta2), p(thr_bad), p(bpm), p(bps), 0, p(out))
    assert rc == 1 and b"C = 30 outside" in L.shim_last_error()
    # cubes with 2 slices through the same entry point, per-session and with one shared curve
    y2, th2, thr2, f2, b2, sds3 = _recover_inputs(n, m, Ccat, 2, 31)
    out2 = np.zeros((g.NGRID, m, 2), order="F")
    for cirf in (0, 1):
        rc = L.shim_recover_fstar(55, u64(n), u64(m), u64(2), u64(Ccat), p(f2), p(y2), p(th2), p(b2), p(thr2), p(bpm), p(bps), cirf, p(out2))
        assert rc == 0, L.shim_last_error().decode()
        assert np.array_equal(out2, g.recover_fstar(55, f2, y2, th2, b2, thr2, beta_prior_sds=bps, constant_IRF=cirf)["fstar"])
        assert np.array_equal(out2[:, :, 0], out2[:, :, 1]) == bool(cirf)


def test_r_shim_gpirtMCMC_at_horizon_2():
    """The shim hands cubes with H slices through unchanged: dynamic GPIRT (GP over time) through the R entry point."""
    import ctypes as C

    import gpirt_b200 as g

    L = _rshim()
    n, m, Ccat, H, sample = 40, 7, 3, 2, 3
    Ps = [make_problem(n, m, Ccat, seed=70 + h) for h in range(H)]
    y = np.asfortranarray(np.stack([P["y"] for P in Ps], axis=2))
    th0 = np.asfortranarray(np.stack([P["theta0"] for P in Ps], axis=1))
    thr = np.asfortranarray(np.stack([Ps[0]["thr0"]] * H, axis=2))
    rng = np.random.default_rng(3)
    z = rng.standard_normal(H * (3 * m + n * m))
    bpm, bps, tpm, tps = np.zeros((3, m), order="F"), Ps[0]["beta_prior_sds"], np.zeros((2, n), order="F"), np.zeros((2, n), order="F")
    out = dict(theta=np.zeros((sample, n, H), order="F"), beta=np.zeros((3, m, H, sample), order="F"), thr=np.zeros((m, Ccat + 1, H, sample), order="F"),
               ll=np.zeros(sample))
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    u64 = C.c_uint64
    rc = L.shim_gpirtMCMC(u64(n), u64(m), u64(H), u64(Ccat), p(y), p(th0), sample, 0, 1, p(bpm), p(bps), p(tpm), p(tps), C.c_double(1.0), C.c_double(1.5),
                          b"RBF", p(thr), 0, p(z), u64(z.size), C.c_double(0.25), p(out["theta"]), p(out["beta"]), p(out["thr"]), p(out["ll"]), None, None)
    assert rc == 0, L.shim_last_error().decode()
    beta0 = np.asfortranarray(np.stack([bpm + bps * z[h * 3 * m:(h + 1) * 3 * m].reshape(m, 3).T for h in range(H)], axis=2))
    z0 = np.asfortranarray(z[H * 3 * m:].reshape(H, m, n).transpose(2, 1, 0))
    ref = g.gpirtMCMC(y, sample, 0, THIN=1, CHAIN=1, beta_prior_means=bpm, beta_prior_sds=bps, theta_prior_sds=tps, theta_os=1.0, theta_ls=1.5, KERNEL="RBF",
                      theta_init=th0, thresholds=thr, SEED=int(0.25 * 4294967296.0), beta_init=beta0, f_init_z=z0)[0]
    assert np.array_equal(out["theta"], ref["theta"]) and np.array_equal(out["ll"], ref["ll"])
    for s in range(sample):
        assert np.array_equal(out["beta"][..., s], ref["beta"][s]) and np.array_equal(out["thr"][..., s], ref["threshold"][s])


def test_r_shim_gpirtMCMC_with_constant_IRF():
    """constant_IRF = 1 through the R entry point: the shim draws beta for session 0 and one N(0,1) vector per item with R's RNG — 3m + nm draws,
    the reference's constant branch (src/gpirtMCMC.cpp:176-188) — and the library composes the mode (mcmc_constant); thresholds are shared by the
    sessions in every kept draw, and the result equals gpirtMCMC() given the same initial values."""
    import ctypes as C

    import gpirt_b200 as g

    L = _rshim()
    n, m, Ccat, H, sample = 40, 7, 3, 2, 3
    Ps = [make_problem(n, m, Ccat, seed=170 + h) for h in range(H)]
    y = np.asfortranarray(np.stack([P["y"] for P in Ps], axis=2))
    th0 = np.asfortranarray(np.stack([P["theta0"] for P in Ps], axis=1))
    thr = np.asfortranarray(np.stack([Ps[0]["thr0"]] * H, axis=2))
    z = np.random.default_rng(3).standard_normal(3 * m + n * m)  # exactly what the constant branch consumes: a longer stream is not needed
    bpm, bps, tpm, tps = np.zeros((3, m), order="F"), Ps[0]["beta_prior_sds"], np.zeros((2, n), order="F"), np.zeros((2, n), order="F")
    out = dict(theta=np.zeros((sample, n, H), order="F"), beta=np.zeros((3, m, H, sample), order="F"), thr=np.zeros((m, Ccat + 1, H, sample), order="F"),
               ll=np.zeros(sample))
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    u64 = C.c_uint64
    rc = L.shim_gpirtMCMC(u64(n), u64(m), u64(H), u64(Ccat), p(y), p(th0), sample, 0, 1, p(bpm), p(bps), p(tpm), p(tps), C.c_double(1.0), C.c_double(1.5),
                          b"Matern", p(thr), 1, p(z), u64(z.size), C.c_double(0.25), p(out["theta"]), p(out["beta"]), p(out["thr"]), p(out["ll"]), None, None)
    assert rc == 0, L.shim_last_error().decode()
    beta0 = np.zeros((3, m, H), order="F")
    beta0[:, :, 0] = bpm + bps * z[:3 * m].reshape(m, 3).T
    z0 = np.zeros((n, m, H), order="F")
    z0[:, :, 0] = z[3 * m:].reshape(m, n).T
    ref = g.gpirtMCMC(y, sample, 0, THIN=1, CHAIN=1, beta_prior_means=bpm, beta_prior_sds=bps, theta_prior_sds=tps, theta_os=1.0, theta_ls=1.5, KERNEL="Matern",
                      theta_init=th0, thresholds=thr, SEED=int(0.25 * 4294967296.0), beta_init=beta0, f_init_z=z0, constant_IRF=1)[0]
    assert np.array_equal(out["theta"], ref["theta"]) and np.array_equal(out["ll"], ref["ll"]) and np.all(np.isfinite(out["ll"]))
    for s in range(sample):
        assert np.array_equal(out["beta"][..., s], ref["beta"][s]) and np.array_equal(out["thr"][..., s], ref["threshold"][s])
        assert np.array_equal(out["thr"][:, :, 0, s], out["thr"][:, :, 1, s])


@pytest.mark.parametrize("n,m,C,chains", [(150, 40, 5, 4), (100, 418, 2, 3), (500, 64, 2, 8)])
def test_batched_chains_equal_the_chains_run_one_at_a_time(n, m, C, chains):
    """BASELINE config 5 (independent chains batched per GPU): gpirtMCMC(CHAIN = k) runs the k chains side by side in one gpirt_mcmc_chains
    call (shared y on the device, round-robin sweeps on per-chain streams). Every chain's kept draws are bit-identical to the same chain
    run alone through gpirt_mcmc — the batching is scheduling only — and the chains differ from each other."""
    import gpirt_b200 as g

    P = make_problem(n, m, C, seed=77 + n)
    kw = dict(beta_prior_sds=P["beta_prior_sds"], theta_prior_sds=P["theta_prior_sds"], theta_init=P["theta0"], thresholds=P["thr0"], SEED=5,
              store_f=True, store_fstar=True)
    together = g.gpirtMCMC(P["y"], 6, 2, THIN=2, CHAIN=chains, **kw)
    assert len(together) == chains
    for c in range(chains):
        alone = g.gpirtMCMC(P["y"], 6, 2, THIN=2, CHAIN=1, first_chain=c, **kw)[0]
        for key in ("theta", "ll"):
            assert np.array_equal(together[c][key], alone[key]), f"chain {c}: {key} differs between the batched and the single run"
        for key in ("beta", "threshold", "f", "fstar"):
            for s in range(3):
                assert np.array_equal(together[c][key][s], alone[key][s]), f"chain {c}: {key} draw {s} differs"
    assert not np.array_equal(together[0]["theta"], together[1]["theta"])


@pytest.mark.parametrize("screen", [1, 0])
def test_a_NaN_state_stalls_the_ESS_loop_instead_of_scoring_finite_garbage(screen):
    """A NaN in the latent state makes every likelihood NaN: std::log / R::pnorm carry it in the reference, where `ll > log_y` is then never
    true and the shrink loop spins for ever (src/draw-f.cpp:23-41 has no cap). Here the fp32 screen must not decide such a proposal (its bound
    becomes NaN) and the fp64 pass must not turn NaN into a finite number (log_pos reads exponent bits): the loop runs into
    ESS_MAX_PROPOSALS and the call returns GPIRT_ERR_ESS_STALLED."""
    import gpirt_b200
    from gpirt_b200 import _lib

    n, m, C = 96, 6, 4
    P = make_problem(n, m, C, seed=3)
    S = gpirt_b200.Sampler(n, m, C, seed=11)
    S.set_screen(screen)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    f0 = np.zeros((n, m), order="F")
    f0[7, 2] = np.nan
    P["y"][7, 2] = 2.0  # the poisoned cell is observed
    S.set_data(P["y"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], f=f0, thresholds=P["thr0"])
    rng = np.random.default_rng(0)
    nu = np.asfortranarray(rng.standard_normal((n, m)))
    with pytest.raises(gpirt_b200.GpirtError) as e:
        S.draw_f(nu=nu)
        S.sync()
    assert e.value.code == _lib.ERR_ESS_STALLED
    S.close()


@pytest.mark.parametrize("shards", [2, 3])
def test_gpirtMCMC_item_sharded_from_one_process_returns_the_single_device_draws(shards, monkeypatch):
    """gpirt_mcmc(n_devices > 1): the chain item-sharded over several shards from ONE process, exchanges as device / peer copies (csrc/entry.cu
    mcmc_sharded; the north star's item partition behind the Rcpp entry point, src/RcppExports.cpp:16-39). Here every shard sits on device 0 —
    the same code path as distinct GPUs, peer copies included — and must return the draws of the one-device call bit for bit; the
    GPIRT_DEVICES environment variable reaches the same path for a caller that cannot change its arguments (the R shim)."""
    import gpirt_b200

    n, m, C = 180, 67, 4
    P = make_problem(n, m, C, seed=17, missing=0.08)
    kw = dict(beta_prior_sds=P["beta_prior_sds"], theta_prior_sds=P["theta_prior_sds"], theta_init=P["theta0"], thresholds=P["thr0"], SEED=31)
    one = gpirt_b200.gpirtMCMC(P["y"], 6, 2, THIN=2, **kw)[0]
    many = gpirt_b200.gpirtMCMC(P["y"], 6, 2, THIN=2, devices=[0] * shards, **kw)[0]
    monkeypatch.setenv("GPIRT_DEVICES", ",".join(["0"] * shards))
    env = gpirt_b200.gpirtMCMC(P["y"], 6, 2, THIN=2, **kw)[0]
    monkeypatch.delenv("GPIRT_DEVICES")
    for other in (many, env):
        assert np.array_equal(one["theta"], other["theta"]) and np.array_equal(one["ll"], other["ll"])
        assert len(other["beta"]) == 3
        for a, b in zip(one["beta"] + one["threshold"], other["beta"] + other["threshold"]):
            assert np.array_equal(a, b)
    with pytest.raises(gpirt_b200.GpirtError) as e:  # f draws live on the shard that owns the item
        gpirt_b200.gpirtMCMC(P["y"], 2, 0, devices=[0, 0], store_f=True, **kw)
    assert e.value.code == 6
