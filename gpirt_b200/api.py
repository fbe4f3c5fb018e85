"""The reference's two entry points and its numerical leaves, over the C-ABI.

gpirtMCMC()/recover_fstar() take the arguments of R/gpirtMCMC.R:98-109 and R/recover_fstar.R:36-39 and
return what the Rcpp functions return (src/gpirtMCMC.cpp:396-403, src/recover_fstar.cpp:92): a list with
one dict per chain holding theta (n_samples, n, 1), f, beta, fstar, threshold (lists) and ll.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import NGRID, GpirtError, McmcArgs, RecoverArgs, check
from .sampler import _F, _p, default_thresholds


def _as_cube(data):
    y = np.asarray(data, dtype=np.float64)
    if y.ndim == 2:
        y = y[:, :, None]
    if y.ndim != 3:
        raise GpirtError(_lib.ERR_SHAPE, "data must be an n x m x horizon array of codes 1..C (NA = missing)")
    return np.asfortranarray(y)


def gpirtMCMC(data, sample_iterations, burn_iterations, THIN=1, CHAIN=1, vote_codes=None, beta_prior_means=None,
              beta_prior_sds=None, theta_prior_means=None, theta_prior_sds=None, theta_os=1, theta_ls=10, KERNEL="Matern",
              theta_init=None, thresholds=None, SEED=1, constant_IRF=0, store_f=False, store_fstar=False, beta_init=None,
              f_init=None, device=0, progress=None, f_init_z=None, first_chain=0, devices=None, verbose=False):
    """Drop-in for gpirtMCMC() (R/gpirtMCMC.R:98-167). `vote_codes` must stay None: the reference's recoding
    path is broken for the ordinal sampler; pass integer codes 1..C with NaN for missing.
    `devices=[0, 1, ...]` (or the environment variable GPIRT_DEVICES="0,1,...") shards the items of every chain over several GPUs."""
    if vote_codes is not None:
        raise GpirtError(_lib.ERR_UNSUPPORTED, "vote_codes recoding is not part of the sampler path; pass codes 1..C")
    y = _as_cube(data)
    n, m, H = y.shape
    bpm = _F(np.zeros((3, m)) if beta_prior_means is None else beta_prior_means)
    bps = _F(np.full((3, m), 3.0) if beta_prior_sds is None else beta_prior_sds)
    tpm = _F(np.zeros((2, n)) if theta_prior_means is None else theta_prior_means)
    tps = _F(np.zeros((2, n)) if theta_prior_sds is None else theta_prior_sds)
    for name, a, shp in (("beta_prior_means", bpm, (3, m)), ("beta_prior_sds", bps, (3, m)), ("theta_prior_means", tpm, (2, n)),
                         ("theta_prior_sds", tps, (2, n))):
        if a.shape != shp:
            raise GpirtError(_lib.ERR_SHAPE, f"{name} has shape {a.shape}, expected {shp}")
    if thresholds is None:
        codes = np.unique(y[np.isfinite(y)])
        Ccat = len(codes)  # R/gpirtMCMC.R:145
        thr = default_thresholds(m, Ccat)[:, :, None].repeat(H, axis=2)
    else:
        thr = np.asarray(thresholds, dtype=np.float64)
        if thr.ndim == 2:  # one m x (C+1) matrix: the same initial thresholds in every session
            thr = thr[:, :, None].repeat(H, axis=2)
        if thr.shape != (m, thr.shape[1], H):
            raise GpirtError(_lib.ERR_SHAPE, f"thresholds has shape {thr.shape}, expected {(m, thr.shape[1], H)}")
        Ccat = thr.shape[1] - 1
    thr = np.asfortranarray(thr)
    n_samples = int(sample_iterations / THIN)
    outs, keep = [], []
    args = (McmcArgs * CHAIN)()
    cb = _lib.PROGRESS_FN(progress) if progress is not None else _lib.PROGRESS_FN()
    # devices=[0, 1, ...]: every chain item-sharded over these GPUs from this one process (gpirt_mcmc_args.n_devices; no torch, no NCCL)
    devs = None if devices is None else np.ascontiguousarray(np.asarray(devices, dtype=np.int32))
    for chain in range(CHAIN):
        rng = np.random.default_rng(SEED + first_chain + chain)  # stands in for set.seed(SEED+chain-1) (R/gpirtMCMC.R:114)
        if theta_init is None:
            th0 = np.empty((n, H), order="F")
            th0[:, 0] = tpm[0] + tps[0] * rng.standard_normal(n)  # rnorm(1, mean, sd) per respondent (:124-126)
            th0[:, 1:] = th0[:, :1]
            theta_init = th0  # as in R: the first chain's draw is assigned in the function's scope, later chains find it non-NULL (:120-134)
        else:
            th0 = _F(np.asarray(theta_init, dtype=np.float64).reshape(n, H))
        out = dict(theta=np.zeros((n_samples, n, H), order="F"), beta=np.zeros((3, m, H, n_samples), order="F"),
                   threshold=np.zeros((m, Ccat + 1, H, n_samples), order="F"), ll=np.zeros(n_samples),
                   f=np.zeros((n, m, H, n_samples), order="F") if store_f else None,
                   fstar=np.zeros((NGRID, m, H, n_samples), order="F") if store_fstar else None)
        a = args[chain]
        b0, f0, z0 = _F(beta_init), _F(f_init), _F(f_init_z)
        keep.append((th0, b0, f0, z0))
        a.y, a.n, a.m, a.horizon = y.ctypes.data, n, m, H
        a.theta = th0.ctypes.data
        a.sample_iterations, a.burn_iterations, a.THIN = int(sample_iterations), int(burn_iterations), int(THIN)
        a.beta_prior_means, a.beta_prior_sds = bpm.ctypes.data, bps.ctypes.data
        a.theta_prior_means, a.theta_prior_sds = tpm.ctypes.data, tps.ctypes.data
        a.theta_os, a.theta_ls, a.KERNEL = float(theta_os), float(theta_ls), KERNEL.encode()
        a.thresholds, a.C = thr.ctypes.data, Ccat
        a.constant_IRF, a.store_f, a.store_fstar = int(constant_IRF), int(store_f), int(store_fstar)
        a.beta_init = b0.ctypes.data if b0 is not None else None
        a.f_init = f0.ctypes.data if f0 is not None else None
        a.f_init_z = z0.ctypes.data if z0 is not None else None  # f0_j = L z_j on the device (rmvnorm(L), src/mvnormal.h:7-14)
        a.seed, a.chain, a.device = int(SEED), first_chain + chain, int(device)
        a.theta_draws, a.beta_draws = out["theta"].ctypes.data, out["beta"].ctypes.data
        a.threshold_draws, a.ll_draws = out["threshold"].ctypes.data, out["ll"].ctypes.data
        a.f_draws = out["f"].ctypes.data if store_f else None
        a.fstar_draws = out["fstar"].ctypes.data if store_fstar else None
        a.progress = cb
        if devs is not None and devs.size > 1:
            a.n_devices, a.devices = int(devs.size), devs.ctypes.data
        outs.append(out)
    if verbose:  # the block the reference prints before the chain starts (src/gpirtMCMC.cpp:60-82)
        print(memory_estimate(args[0])[1], end="")
    # all chains in one call: they share y on the device and run side by side (gpirt_mcmc_chains; CHAIN == 1 is gpirt_mcmc)
    check(_lib.lib().gpirt_mcmc_chains(args, CHAIN) if CHAIN > 1 else _lib.lib().gpirt_mcmc(C.byref(args[0])))
    result = []
    for out in outs:
        result.append(dict(theta=out["theta"],
                           f=[out["f"][..., s] for s in range(n_samples)] if store_f else None,
                           beta=[out["beta"][..., s] for s in range(n_samples)],
                           fstar=[out["fstar"][..., s] for s in range(n_samples)] if store_fstar else None,
                           threshold=[out["threshold"][..., s] for s in range(n_samples)], ll=out["ll"]))
    return result


def _gpirtMCMC_constant_irf(y, sample_iterations, burn_iterations, THIN, CHAIN, beta_prior_means, beta_prior_sds, theta_prior_means,
                            theta_prior_sds, theta_os, theta_ls, KERNEL, theta_init, thresholds, SEED, store_f, store_fstar, device, progress,
                            first_chain):
    """gpirtMCMC(constant_IRF = 1) step by step: the sweep loop of src/gpirtMCMC.cpp:261-356 over a ConstantPanel (constant_panel.py). The C entry
    point gpirt_mcmc runs the same composition (csrc/entry.cu mcmc_constant) and is what gpirtMCMC() calls; this host loop exists so the tests
    can hold the entry point to the block-by-block chain, bit for bit."""
    from .constant_panel import ConstantPanel

    n, m, H = y.shape
    bpm = _F(np.zeros((3, m)) if beta_prior_means is None else beta_prior_means)
    bps = _F(np.full((3, m), 3.0) if beta_prior_sds is None else beta_prior_sds)
    tpm = _F(np.zeros((2, n)) if theta_prior_means is None else theta_prior_means)
    tps = _F(np.zeros((2, n)) if theta_prior_sds is None else theta_prior_sds)
    if thresholds is None:
        Ccat = len(np.unique(y[np.isfinite(y)]))
        thr = default_thresholds(m, Ccat)[:, :, None].repeat(H, axis=2)
    else:
        thr = np.asarray(thresholds, dtype=np.float64)
        thr = thr[:, :, None].repeat(H, axis=2) if thr.ndim == 2 else thr
        Ccat = thr.shape[1] - 1
    n_samples = int(sample_iterations / THIN)
    result = []
    for chain in range(CHAIN):
        rng = np.random.default_rng(SEED + first_chain + chain)
        if theta_init is None:
            th0 = np.empty((n, H), order="F")
            th0[:, 0] = tpm[0] + tps[0] * rng.standard_normal(n)
            th0[:, 1:] = th0[:, :1]
            theta_init = th0
        th0 = np.clip(np.asarray(theta_init, dtype=np.float64).reshape(n, H), -5.0, 5.0)
        P = ConstantPanel(n, m, Ccat, H, theta_os=theta_os, theta_ls=theta_ls, KERNEL=KERNEL, seed=SEED, chain=first_chain + chain, device=device)
        try:
            P.set_data(y)
            P.set_priors(bps, tps)
            # the constant branch of the initialisation (src/gpirtMCMC.cpp:164-200): beta for session 0 only, one f0 = L_0 z per item for all sessions
            S0 = P.S[0]
            S0.set_state(theta=th0[:, 0], thresholds=thr[:, :, 0])
            S0.set_iter(0xFFFFFFFF)
            S0.init_draw(beta_prior_means=bpm, draw_beta=True, draw_f=True)
            st0 = S0.get_state()
            P.init_constant(th0, st0["beta"], st0["f"], thr)
            P.set_iter(0xFFFFFFFF)
            P.draw_fstar()  # src/gpirtMCMC.cpp:227
            P.set_iter(0)
            out = dict(theta=np.zeros((n_samples, n, H), order="F"), beta=[], threshold=[], ll=np.zeros(n_samples), f=[] if store_f else None,
                       fstar=[] if store_fstar else None)
            total = sample_iterations + burn_iterations
            for it in range(total):
                if progress is not None and progress(it, total, None):
                    raise GpirtError(_lib.ERR_INVALID, "interrupted")
                P.sweep()
                if it >= burn_iterations and it % THIN == 0 and (it - burn_iterations) // THIN < n_samples:  # src/gpirtMCMC.cpp:334-335
                    k = (it - burn_iterations) // THIN
                    s = P.get_state(fstar=True)
                    out["theta"][k] = s["theta"]
                    out["beta"].append(s["beta"])
                    out["threshold"].append(s["thresholds"])
                    out["ll"][k] = P.total_ll()
                    if store_f:
                        out["f"].append(s["f"])
                    if store_fstar:
                        out["fstar"].append(s["fstar"])
            result.append(out)
        finally:
            P.close()
    return result


def recover_fstar(seed, f, y, theta, beta, thresholds, beta_prior_means=None, beta_prior_sds=None, constant_IRF=0, device=0):
    """Drop-in for recover_fstar() (R/recover_fstar.R:36-46): one draw_f + draw_fstar from a stored f and a seed."""
    yy = _as_cube(y)
    n, m, H = yy.shape
    ff = np.asfortranarray(np.asarray(f, dtype=np.float64).reshape(n, m, H))
    th = _F(np.asarray(theta, dtype=np.float64).reshape(n, H))
    be = np.asfortranarray(np.asarray(beta, dtype=np.float64).reshape(2, m, H))
    thr = np.asarray(thresholds, dtype=np.float64)
    thr = np.asfortranarray(np.repeat(thr[:, :, None], H, axis=2) if thr.ndim == 2 else thr)
    if thr.shape[0] != m or thr.shape[2] != H:
        raise ValueError("thresholds must be m x (C+1) or m x (C+1) x horizon")
    bpm = _F(np.zeros((2, m)) if beta_prior_means is None else beta_prior_means)
    bps = _F(np.full((2, m), 0.5) if beta_prior_sds is None else beta_prior_sds)
    out = np.zeros((NGRID, m, H), order="F")
    a = RecoverArgs()
    a.seed, a.f, a.y, a.n, a.m, a.horizon = int(seed), ff.ctypes.data, yy.ctypes.data, n, m, H
    a.theta, a.beta, a.thresholds, a.C = th.ctypes.data, be.ctypes.data, thr.ctypes.data, thr.shape[1] - 1
    a.beta_prior_means, a.beta_prior_sds = bpm.ctypes.data, bps.ctypes.data
    a.constant_IRF, a.device, a.fstar = int(constant_IRF), int(device), out.ctypes.data
    check(_lib.lib().gpirt_recover_fstar(C.byref(a)))
    return dict(fstar=out)


# ---- numerical leaves (src/gpirt.h:262-290) ----
def K(x1, x2, beta_prior_sds, device=0):
    x1, x2, s = _F(x1).ravel(), _F(x2).ravel(), _F(beta_prior_sds).ravel()
    out = np.empty((x1.size, x2.size), order="F")
    check(_lib.lib().gpirt_K(device, _p(x1), C.c_int64(x1.size), _p(x2), C.c_int64(x2.size), _p(s), _p(out)))
    return out


def K_time(x1, x2, os, ls, theta_prior_sds, KERNEL, device=0):
    x1, x2, s = _F(x1).ravel(), _F(x2).ravel(), _F(theta_prior_sds).ravel()
    out = np.empty((x1.size, x2.size), order="F")
    kern = {"Matern": 0, "RBF": 1}.get(KERNEL, 2)
    check(_lib.lib().gpirt_K_time(device, _p(x1), C.c_int64(x1.size), _p(x2), C.c_int64(x2.size), C.c_double(os), C.c_double(ls),
                                  _p(s), kern, _p(out)))
    return out


def chol(A, device=0):
    A = _F(A)
    out = np.empty_like(A, order="F")
    check(_lib.lib().gpirt_chol(device, _p(A), C.c_int64(A.shape[0]), _p(out)))
    return out


def double_solve(L, X, device=0):
    L = _F(L)
    X = _F(np.asarray(X, dtype=np.float64).reshape(L.shape[0], -1))
    out = np.empty_like(X, order="F")
    check(_lib.lib().gpirt_double_solve(device, _p(L), C.c_int64(L.shape[0]), _p(X), C.c_int64(X.shape[1]), _p(out)))
    return out


def interp1(X, Y, XI, device=0):
    """arma::interp1(X, Y[:, j], XI, "linear") for every column j, on the GPU (gpirt_interp1; src/draw-fstar.cpp:77-82)."""
    X, Y, XI = _F(np.asarray(X, dtype=np.float64).ravel()), _F(np.asarray(Y, dtype=np.float64)), _F(np.asarray(XI, dtype=np.float64).ravel())
    if Y.ndim == 1:
        Y = _F(Y[:, None])
    out = np.empty((XI.size, Y.shape[1]), order="F")
    check(_lib.lib().gpirt_interp1(int(device), _p(X), C.c_int64(X.size), _p(Y), C.c_int64(Y.shape[1]), _p(XI), C.c_int64(XI.size), _p(out)))
    return out


def delta_to_threshold(deltas):
    d = _F(deltas).ravel()
    out = np.empty(d.size + 2)
    check(_lib.lib().gpirt_delta_to_threshold(_p(d), d.size + 1, _p(out)))
    return out


def threshold_to_delta(thresholds):
    t = _F(thresholds).ravel()
    out = np.empty(t.size - 2)
    check(_lib.lib().gpirt_threshold_to_delta(_p(t), t.size - 1, _p(out)))
    return out


def debug_probit(x, device=0):
    """(Phi(x), log(x)) as the likelihood kernels evaluate them on the device (test hook)."""
    x = _F(x).ravel()
    ph, lg = np.empty_like(x), np.empty_like(x)
    check(_lib.lib().gpirt_debug_probit(device, _p(x), int(x.size), _p(ph), _p(lg)))
    return ph, lg


def debug_cell32(g, tau_lo, tau_hi, device=0):
    """(t32, e32, t64): the fp32 screening term, its error bound and the fp64 cell term, per element (test hook)."""
    g, lo, hi = _F(g).ravel(), _F(tau_lo).ravel(), _F(tau_hi).ravel()
    t32, e32, t64 = np.empty_like(g), np.empty_like(g), np.empty_like(g)
    check(_lib.lib().gpirt_debug_cell32(device, _p(g), _p(lo), _p(hi), int(g.size), _p(t32), _p(e32), _p(t64)))
    return t32, e32, t64


def fp64_peak(device=0):
    """(DMMA.8x8x4, DFMA) issue-rate peaks of the device in TFLOP/s, measured now (bench.py's fp64 roofline denominator)."""
    import ctypes as C

    a, b = C.c_double(), C.c_double()
    check(_lib.lib().gpirt_debug_fp64_peak(int(device), C.byref(a), C.byref(b)))
    return a.value, b.value


def memory_estimate(args):
    """(numbers, text) of gpirtMCMC()'s memory estimate (src/gpirtMCMC.cpp:47-82) for a filled McmcArgs; text is the reference's printed block."""
    est = _lib.MemEstimate()
    buf = C.create_string_buffer(2048)
    check(_lib.lib().gpirt_memory_estimate(C.byref(args), C.byref(est), buf, C.c_int64(len(buf))))
    return {k: getattr(est, k) for k, _ in est._fields_}, buf.value.decode()


def debug_phi_tab(x, device=0):
    """phi_tab(x) element-wise on the device: the table-based Phi of k_item_ll / k_theta_table"""
    x = np.ascontiguousarray(np.asarray(x, dtype=np.float64))
    out = np.empty_like(x)
    check(_lib.lib().gpirt_debug_phi_tab(int(device), _p(x), int(x.size), _p(out)))
    return out
