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
