"""Generates tests/golden/*.npz from the REFERENCE'S OWN CODE (oracle/_ref/libgpirt_ref.so = /root/reference/src/*.cpp compiled
against oracle/refshim). Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py

Each sweep case stores
  * the inputs of gpirtMCMC() (src/gpirtMCMC.cpp:25-38) and the N(0,1) stream its R::rnorm initialisation consumed,
  * the reference's stored draws after every sweep (theta, beta, threshold, ll, f, fstar), single OpenMP thread,
  * the random streams each unit consumed in each block of each sweep. The reference draws them from per-thread mt19937
    generators (src/gpirt.h:13-41,178-182); they are captured by running the CPU oracle in its mt19937 mode beside the reference
    and are only kept if the oracle reproduces the reference's theta draws bit for bit (asserted below), i.e. the recorded streams
    ARE the reference's. The CUDA path and the oracle replay them as supplied streams (tests/test_golden.py).
The leaf file stores direct outputs of the reference's K, K_time, ll_bar_sparse, threshold transforms, update_cholesky_cache and
double_solve on small inputs.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import orc, ref  # noqa: E402
from gpirt_b200 import response_matrix as rmx  # noqa: E402
from gpirt_b200.sampler import default_thresholds  # noqa: E402
from tests.problems import make_problem  # noqa: E402

UCAP = 48


def sweep_case(name, n, m, C, seed, iters, missing, empty_item=None, tps_sd=0.0):
    P = make_problem(n, m, C, seed=seed, missing=missing)
    y = P["y"].copy(order="F")
    if empty_item is not None:
        y[:, empty_item] = np.nan  # an item nobody answered: draw_beta keeps its beta and draws nothing (src/draw-beta.cpp:73,97-99)
    rng = np.random.default_rng(seed + 1000)
    z_init = rng.standard_normal(3 * m + n * m)
    bpm = np.zeros((3, m), order="F")
    bps = P["beta_prior_sds"]
    tpm = np.zeros((2, n), order="F")
    tps = np.full((2, n), tps_sd, order="F")
    theta_in = P["theta0"] * 1.7  # some values beyond +-5 so the clamp (src/gpirtMCMC.cpp:133) is exercised
    R = ref.gpirtMCMC(y, theta_in, iters, 0, 1, bpm, bps, tpm, tps, 1.0, 10.0, "Matern", P["thr0"], 0, z_init, store_f=True, store_fstar=True)

    # the oracle beside it, mt19937 mode, recording every unit's draws
    theta0 = np.clip(theta_in, -5.0, 5.0)
    beta0 = np.asfortranarray(bpm + bps * z_init[:3 * m].reshape(m, 3).T)  # R::rnorm(mean, sd) per (item, coefficient)
    L0 = ref.update_cholesky_cache(theta0, bps)[0][:, :, 0]
    f0 = np.asfortranarray(np.stack([L0 @ z_init[3 * m + j * n:3 * m + (j + 1) * n] for j in range(m)], axis=1))
    O = orc.Sampler(y, C, bps, tps)
    O.set_state(theta=theta0, beta=beta0, f=f0, thr=P["thr0"])
    O.compute_mu()
    assert O.update_cholesky() == 0
    O.draw_fstar(orc.mt())  # the initial draw_fstar (src/gpirtMCMC.cpp:227) runs on the unseeded generators
    fstar0 = O.state()["fstar"]
    O.record(True)
    rec = {}
    for it in range(iters):
        O.mt_seed_all(it * 10000)
        assert O.sweep(orc.mt()) == 0
        for blk, zcap in (("f", n), ("fstar", orc.NGRID), ("theta", 1), ("beta", 3), ("thr", C - 1)):
            z, u, nz, nu = O.recorded(blk, zcap, UCAP)
            rec[f"z_{blk}_{it}"] = z
            if blk != "fstar":
                rec[f"u_{blk}_{it}"] = u
                rec[f"nu_{blk}_{it}"] = nu
        s = O.state()
        assert np.array_equal(s["theta"], R["theta"][it, :, 0]), f"{name}: oracle theta != reference theta at sweep {it}"
        assert np.array_equal(s["beta"], R["beta"][:, :, 0, it]) and np.array_equal(s["thr"][:, 1:-1], R["threshold"][:, 1:-1, 0, it])
    out = dict(n=n, m=m, C=C, iters=iters, y=y, theta_in=theta_in, z_init=z_init, beta_prior_means=bpm, beta_prior_sds=bps,
               theta_prior_sds=tps, thr0=P["thr0"], beta0=beta0, f0=f0, fstar0=fstar0,
               ref_theta=R["theta"][:, :, 0], ref_beta=R["beta"][:, :, 0, :], ref_threshold=R["threshold"][:, :, 0, :], ref_ll=R["ll"],
               ref_f=R["f"][:, :, 0, :], ref_fstar_last=R["fstar"][:, :, 0, -1], **rec)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "ll", R["ll"], "max evals recorded", max(int(v.max()) for k, v in rec.items() if k.startswith("nu_")))


def panel_case(name, n, m, C, H, seed, iters, ls, kernel, os_=1.3):
    """horizon > 1: the reference's gpirtMCMC at H sessions (RDM for ls <= 0.1, GP over time for 0.1 < ls < 3H) and the streams every unit of
    every session consumed, captured from the oracle's panel in mt19937 mode run beside it (kept only if it reproduces the reference's theta,
    beta and thresholds bit for bit). Keys: z_<block>_<it>_<h> / u_... / nu_... per session; the GP theta step has no session suffix."""
    Ps = [make_problem(n, m, C, seed=seed + h, missing=0.12) for h in range(H)]
    y = np.asfortranarray(np.stack([P["y"] for P in Ps], axis=2))
    theta_in = np.asfortranarray(np.stack([P["theta0"] for P in Ps], axis=1)) * 1.5
    rng = np.random.default_rng(seed + 1000)
    thr0 = np.asfortranarray(np.stack([Ps[0]["thr0"]] * H, axis=2))
    z_init = rng.standard_normal(H * (3 * m + n * m))
    bpm, bps, tpm = np.zeros((3, m), order="F"), Ps[0]["beta_prior_sds"], np.zeros((2, n), order="F")
    tps = np.asfortranarray(np.vstack([rng.uniform(0.0, 0.7, (1, n)), np.zeros((1, n))]))
    R = ref.gpirtMCMC(y, theta_in, iters, 0, 1, bpm, bps, tpm, tps, os_, ls, kernel, thr0, 0, z_init, store_f=True, store_fstar=True)
    theta0 = np.clip(theta_in, -5.0, 5.0)
    beta0 = np.asfortranarray(np.stack([bpm + bps * z_init[h * 3 * m:(h + 1) * 3 * m].reshape(m, 3).T for h in range(H)], axis=2))
    zf = z_init[H * 3 * m:]
    L0 = ref.update_cholesky_cache(theta0, bps)[0]
    f0 = np.asfortranarray(np.stack([np.stack([L0[:, :, h] @ zf[(h * m + j) * n:(h * m + j + 1) * n] for j in range(m)], axis=1) for h in range(H)], axis=2))
    O = orc.Panel(y, C, bps, tps, theta_os=os_, theta_ls=ls, KERNEL=kernel)
    O.set_state(theta=theta0, beta=beta0, f=f0, thr=thr0)
    O.compute_mu()
    assert O.update_cholesky() == 0
    for S in O.S:
        S.draw_fstar(orc.mt())  # the initial draw_fstar (src/gpirtMCMC.cpp:227), unseeded generators
    fstar0 = O.state()["fstar"]
    O.record(True)
    rec = {}
    for it in range(iters):
        O.mt_seed_all(it * 10000)
        assert O.sweep(lambda block, h: orc.mt()) == 0
        for h, S in enumerate(O.S):
            blocks = [("f", n), ("fstar", orc.NGRID), ("beta", 3), ("thr", C - 1)] + ([("theta", 1)] if O.regime == 1 else [])
            for blk, zcap in blocks:
                z, u, nz, nu = S.recorded(blk, zcap, UCAP)
                rec[f"z_{blk}_{it}_{h}"] = z
                if blk != "fstar":
                    rec[f"u_{blk}_{it}_{h}"] = u
                    rec[f"nu_{blk}_{it}_{h}"] = nu
        if O.regime == 2:
            z, u, nz, nu = O.S[0].recorded("theta", H, UCAP)
            rec[f"z_theta_{it}"], rec[f"u_theta_{it}"], rec[f"nu_theta_{it}"] = z, u, nu
        s = O.state()
        assert np.array_equal(s["theta"], R["theta"][it]), f"{name}: oracle theta != reference theta at sweep {it}"
        assert np.array_equal(s["beta"], R["beta"][..., it]) and np.array_equal(s["thr"][:, 1:-1], R["threshold"][:, 1:-1, :, it])
    out = dict(n=n, m=m, C=C, H=H, iters=iters, ls=ls, os=os_, kernel=kernel, y=y, theta_in=theta_in, beta_prior_sds=bps, theta_prior_sds=tps,
               thr0=thr0, beta0=beta0, f0=f0, fstar0=fstar0, ref_theta=R["theta"], ref_beta=R["beta"], ref_threshold=R["threshold"], ref_ll=R["ll"],
               ref_f=R["f"], ref_fstar_last=R["fstar"][..., -1], **rec)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "ll", R["ll"])


def dataset_case(name, cube, C, seed, iters, keep_items=8):
    """BASELINE configs 1 and 2: the reference's gpirtMCMC() on its own datasets (data/senate116.rda, data/SDO.rda read by
    gpirt_b200.response_matrix) with the R wrapper's defaults (R/gpirtMCMC.R:98-109: beta sds 3, theta_prior_sds 0, os 1, ls 10, Matern,
    thresholds qnorm(i/C)) and theta_init ~ N(0,1) (the wrapper's own default rnorm(mean, sd = 0) starts every respondent at 0).
    The fixture keeps the response cube (int8, 0 = missing), the initialisation stream and the reference's theta / beta / threshold / ll
    draws, f and f* for the first `keep_items` items. The per-unit random streams are NOT stored (10 MB): the tests regenerate them by
    running the oracle in its mt19937 mode, which must reproduce the reference's theta bit for bit before its streams are trusted."""
    y = np.asfortranarray(cube[:, :, 0])
    n, m = y.shape
    rng = np.random.default_rng(seed)
    theta_in = rng.standard_normal(n)
    z_init = rng.standard_normal(3 * m + n * m)
    bpm, bps = np.zeros((3, m), order="F"), np.full((3, m), 3.0, order="F")
    tpm, tps = np.zeros((2, n), order="F"), np.zeros((2, n), order="F")
    thr0 = default_thresholds(m, C)
    R = ref.gpirtMCMC(y, theta_in, iters, 0, 1, bpm, bps, tpm, tps, 1.0, 10.0, "Matern", thr0, 0, z_init, store_f=True, store_fstar=True)
    k = min(keep_items, m)
    out = dict(n=n, m=m, C=C, iters=iters, y=np.where(np.isfinite(y), y, 0).astype(np.int8), theta_in=theta_in, z_init=z_init,
               ref_theta=R["theta"][:, :, 0], ref_beta=R["beta"][:, :, 0, :], ref_threshold=R["threshold"][:, :, 0, :], ref_ll=R["ll"],
               ref_f_head=R["f"][:, :k, 0, :], ref_fstar_last_head=R["fstar"][:, :k, 0, -1])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, y.shape, "C", C, "ll", R["ll"])


def cst_case(name, n, m, C, H, seed, iters, ls):
    """constant_IRF == 1: the reference's own gpirtMCMC run and the streams every unit consumed, captured from the oracle's ConstantPanel in
    mt19937 mode beside it (kept only if it reproduces the reference's theta, beta and thresholds bit for bit). Stacked blocks (f, fstar, thr)
    have one stream set per sweep (units = items over the n*H stacked rows), beta one per session, theta per session (RDM) or one (GP)."""
    Ps = [make_problem(n, m, C, seed=seed + h, missing=0.1) for h in range(H)]
    y = np.asfortranarray(np.stack([P["y"] for P in Ps], axis=2))
    theta_in = np.asfortranarray(np.stack([P["theta0"] for P in Ps], axis=1))
    thr0 = np.asfortranarray(np.stack([Ps[0]["thr0"]] * H, axis=2))
    z_init = np.random.default_rng(seed + 1000).standard_normal(3 * m + n * m)
    bpm, bps, tpm, tps = np.zeros((3, m), order="F"), Ps[0]["beta_prior_sds"], np.zeros((2, n), order="F"), np.zeros((2, n), order="F")
    R = ref.gpirtMCMC(y, theta_in, iters, 0, 1, bpm, bps, tpm, tps, 1.0, ls, "Matern", thr0, 1, z_init, store_f=True, store_fstar=True)
    theta0 = np.clip(theta_in, -5.0, 5.0)
    beta0 = np.asfortranarray(bpm + bps * z_init[:3 * m].reshape(m, 3).T)
    L0 = ref.update_cholesky_cache(theta0, bps)[0][:, :, 0]
    f0 = np.asfortranarray(np.stack([L0 @ z_init[3 * m + j * n:3 * m + (j + 1) * n] for j in range(m)], axis=1))
    O = orc.ConstantPanel(y, C, bps, tps, theta_os=1.0, theta_ls=ls, KERNEL="Matern")
    O.init_constant(theta0, beta0, f0, thr0)
    assert O.draw_fstar(orc.mt()) == 0
    fstar0 = O.state()["fstar"][:, :, 0]
    O.record(True)
    O.T.record(True)
    rec = {}
    for it in range(iters):
        O.mt_seed_all(it * 10000)
        assert O.sweep(lambda block, h: orc.mt()) == 0
        for blk, zcap in (("f", n * H), ("fstar", orc.NGRID), ("thr", C - 1)):
            zz, u, nz, nu = O.T.recorded(blk, zcap, UCAP)
            rec[f"z_{blk}_{it}"] = zz
            if blk != "fstar":
                rec[f"u_{blk}_{it}"], rec[f"nu_{blk}_{it}"] = u, nu
        for h, S in enumerate(O.S):
            zz, u, nz, nu = S.recorded("beta", 3, UCAP)
            rec[f"z_beta_{it}_{h}"], rec[f"u_beta_{it}_{h}"], rec[f"nu_beta_{it}_{h}"] = zz, u, nu
            if O.regime != 2:  # RDM, or one session in the CST regime: the session's own theta streams
                zz, u, nz, nu = S.recorded("theta", 1, UCAP)
                rec[f"z_theta_{it}_{h}"], rec[f"u_theta_{it}_{h}"], rec[f"nu_theta_{it}_{h}"] = zz, u, nu
        if O.regime == 2:
            zz, u, nz, nu = O.S[0].recorded("theta", H, UCAP)
            rec[f"z_theta_{it}"], rec[f"u_theta_{it}"], rec[f"nu_theta_{it}"] = zz, u, nu
        s = O.state()
        assert np.array_equal(s["theta"], R["theta"][it]), f"{name}: oracle theta != reference theta at sweep {it}"
        assert np.array_equal(s["beta"], R["beta"][..., it]) and np.array_equal(s["thr"][:, 1:-1], R["threshold"][:, 1:-1, :, it])
    out = dict(n=n, m=m, C=C, H=H, iters=iters, ls=ls, y=y, theta_in=theta_in, beta_prior_sds=bps, theta_prior_sds=tps, thr0=thr0, beta0=beta0,
               f0=f0, fstar0=fstar0, ref_theta=R["theta"], ref_beta=R["beta"], ref_threshold=R["threshold"], ref_ll=R["ll"], ref_f=R["f"],
               ref_fstar_last=R["fstar"][..., -1], **rec)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "ll", R["ll"])


def leaf_case():
    rng = np.random.default_rng(42)
    x1, x2, sds = rng.standard_normal(17) * 2, rng.standard_normal(9) * 2, np.array([3.0, 2.5, 0.7])
    ts = np.arange(5.0)
    out = dict(x1=x1, x2=x2, sds=sds, K=ref.K(x1, x2, sds), ts=ts, K_time_matern=ref.K_time(ts, ts, 1.3, 2.0, [0.4, 0.2], "Matern"),
               K_time_rbf=ref.K_time(ts, ts, 1.3, 2.0, [0.4, 0.2], "RBF"))
    xs = np.concatenate([np.linspace(-38.5, 9.0, 400), rng.standard_normal(200) * 3, [0.0, -0.67448975, 0.67448975, 5.656854249492381]])
    out.update(pnorm_x=xs, pnorm=ref.pnorm(xs))
    n, C = 40, 5
    f, mu = rng.standard_normal(n) * 2, rng.standard_normal(n)
    y = rng.integers(1, C + 1, n).astype(np.float64)
    y[rng.random(n) < 0.15] = np.nan
    thr = np.array([-np.inf, -1.1, -0.2, 0.5, 1.7, np.inf])
    out.update(ll_f=f, ll_mu=mu, ll_y=y, ll_thr=thr, ll=ref.ll_bar_sparse(f, y, mu, thr), delta=ref.threshold_to_delta(thr),
               thr_back=ref.delta_to_threshold(ref.threshold_to_delta(thr)), thr2=ref.delta_to_threshold([0.25]))
    theta = np.clip(rng.standard_normal(24) * 1.5, -5, 5)
    bps = np.tile(np.array([[0.8], [0.5], [0.1]]), (1, 3))  # well-conditioned: 1e-10 element-wise is meaningful
    L, _ = ref.update_cholesky_cache(theta, bps)
    X = rng.standard_normal((24, 3))
    out.update(chol_theta=theta, chol_sds=bps, chol_L=L[:, :, 0], solve_X=X, solve_out=ref.double_solve(L[:, :, 0], X))
    theta2 = np.tile(theta[:, None], (1, 3))
    _, Lt = ref.update_cholesky_cache(theta2, bps, 1.0, 2.0, "Matern")  # 0.1 < ls < 3H: the K_time branch (src/cholesky-cache.cpp:27-35)
    out.update(L_time=Lt[:, :, 0])
    np.savez_compressed(os.path.join(HERE, "leaves.npz"), **out)
    print("leaves ok")


def datasets():
    cube, C, _, _ = rmx.senate116_responses("/root/reference/data/senate116.rda")
    dataset_case("senate116", cube, C, seed=116, iters=3)
    cube, C = rmx.sdo_responses("/root/reference/data/SDO.rda")
    dataset_case("sdo", cube, C, seed=1500, iters=2)


if __name__ == "__main__":
    assert ref.available(), "build oracle/_ref first (make -C oracle ref)"
    orc.set_threads(1)
    ref.set_threads(1)
    if sys.argv[1:] == ["datasets"]:
        datasets()
        sys.exit(0)
    if sys.argv[1:] == ["cst"]:
        cst_case("cst_rdm_h2", n=30, m=5, C=4, H=2, seed=81, iters=2, ls=0.05)
        cst_case("cst_gp_h2", n=30, m=5, C=3, H=2, seed=91, iters=2, ls=2.5)
        cst_case("cst_cst_h1", n=40, m=6, C=4, H=1, seed=101, iters=2, ls=10.0)
        sys.exit(0)
    leaf_case()
    sweep_case("sweep_ordinal_c4", n=48, m=8, C=4, seed=11, iters=3, missing=0.1)
    sweep_case("sweep_binary_c2", n=36, m=6, C=2, seed=23, iters=3, missing=0.2, empty_item=4, tps_sd=0.6)
    panel_case("panel_rdm_h3", n=30, m=5, C=4, H=3, seed=51, iters=2, ls=0.05, kernel="Matern")
    panel_case("panel_gp_h3", n=30, m=5, C=3, H=3, seed=61, iters=2, ls=2.5, kernel="Matern")
    datasets()
    cst_case("cst_rdm_h2", n=30, m=5, C=4, H=2, seed=81, iters=2, ls=0.05)
    cst_case("cst_gp_h2", n=30, m=5, C=3, H=2, seed=91, iters=2, ls=2.5)
    cst_case("cst_cst_h1", n=40, m=6, C=4, H=1, seed=101, iters=2, ls=10.0)
