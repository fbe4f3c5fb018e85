"""TEST INFRASTRUCTURE — ctypes front-end to oracle/_ref/libgpirt_ref.so: the reference's OWN translation units
(/root/reference/src/*.cpp) compiled against the stand-in headers in oracle/refshim/ (oracle/Makefile `ref`).

Used by tests/test_oracle_vs_ref.py to pin the CPU oracle to the reference's code, and by tests/golden/make_golden.py to
record fixtures. Never imported by gpirt_b200. All matrices column-major float64, y with NaN = missing, horizon H explicit.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libgpirt_ref.so")
NGRID = 1001
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_last_error.restype = C.c_char_p
        L.ref_pnorm.restype = C.c_double
        L.ref_pnorm.argtypes = [C.c_double]
        _lib = L
    return _lib


def _F(a):
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _u(x):
    return C.c_uint64(int(x))


def _ck(rc):
    if rc != 0:
        raise RuntimeError(lib().ref_last_error().decode())


def _cube(a, ndim3=True):
    """n x m or n x m x H -> (column-major n x m x H array, H)"""
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 2:
        a = a[:, :, None]
    return np.asfortranarray(a), a.shape[2]


def set_threads(k):
    lib().ref_set_threads(C.c_int(k))


def pnorm(x):
    return np.array([lib().ref_pnorm(float(v)) for v in np.ravel(x)])


def K(x1, x2, sds):
    x1, x2, sds = _F(x1), _F(x2), _F(sds)
    out = np.empty((x1.size, x2.size), order="F")
    _ck(lib().ref_K(_p(x1), _u(x1.size), _p(x2), _u(x2.size), _p(sds), _p(out)))
    return out


def K_time(x1, x2, os_, ls, tsds, kernel="Matern"):
    x1, x2, tsds = _F(x1), _F(x2), _F(tsds)
    out = np.zeros((x1.size, x2.size), order="F")
    _ck(lib().ref_K_time(_p(x1), _u(x1.size), _p(x2), _u(x2.size), C.c_double(os_), C.c_double(ls), _p(tsds), kernel.encode(), _p(out)))
    return out


def ll_bar_sparse(f, y, mu, thr):
    f, y, mu, thr = _F(f), _F(y), _F(mu), _F(thr)
    out = C.c_double()
    _ck(lib().ref_ll_bar_sparse(_p(f), _p(y), _p(mu), _p(thr), _u(f.size), _u(thr.size - 1), C.byref(out)))
    return out.value


def delta_to_threshold(d):
    d = _F(d)
    Cc = d.size + 1
    out = np.empty(Cc + 1)
    _ck(lib().ref_delta_to_threshold(_p(d), _u(Cc), _p(out)))
    return out


def threshold_to_delta(t):
    t = _F(t)
    Cc = t.size - 1
    out = np.empty(Cc - 1)
    _ck(lib().ref_threshold_to_delta(_p(t), _u(Cc), _p(out)))
    return out


def update_cholesky_cache(theta, sds, os_=0.0, ls=0.0, kernel=""):
    """-> (L n x n x H, L_time H x H x n)"""
    theta = _F(np.asarray(theta, dtype=np.float64).reshape(len(theta), -1))
    n, H = theta.shape
    sds = _F(sds)
    L = np.zeros((n, n, H), order="F")
    Lt = np.zeros((H, H, n), order="F")
    _ck(lib().ref_update_cholesky_cache(_p(theta), _u(n), _u(H), _p(sds), _u(sds.shape[1]), C.c_double(os_), C.c_double(ls), kernel.encode(),
                                        _p(L), _p(Lt)))
    return L, Lt


def double_solve(L, X):
    L = _F(L)
    X = _F(np.asarray(X, dtype=np.float64).reshape(L.shape[0], -1))
    out = np.empty_like(X, order="F")
    _ck(lib().ref_double_solve(_p(L), _u(L.shape[0]), _p(X), _u(X.shape[1]), _p(out)))
    return out


def draw_f(seed, f, theta, y, sds, mu, thr, constant_IRF=0):
    f, H = _cube(f)
    y, _ = _cube(y)
    mu, _ = _cube(mu)
    thr, _ = _cube(thr)
    n, m, _ = f.shape
    theta = _F(np.asarray(theta, dtype=np.float64).reshape(n, H))
    out = np.empty_like(f, order="F")
    _ck(lib().ref_draw_f(C.c_uint32(seed), _u(n), _u(m), _u(H), _u(thr.shape[1] - 1), _p(f), _p(theta), _p(y), _p(_F(sds)), _p(mu), _p(thr),
                         C.c_int(constant_IRF), _p(out)))
    return out


def draw_fstar(seed, f, theta, sds, mu_star, constant_IRF=0):
    f, H = _cube(f)
    mu_star, _ = _cube(mu_star)
    n, m, _ = f.shape
    theta = _F(np.asarray(theta, dtype=np.float64).reshape(n, H))
    out = np.empty((NGRID, m, H), order="F")
    _ck(lib().ref_draw_fstar(C.c_uint32(seed), _u(n), _u(m), _u(H), _p(f), _p(theta), _p(_F(sds)), _p(mu_star), C.c_int(constant_IRF), _p(out)))
    return out


def draw_theta(seed, y, theta, tps, fstar, mu_star, thr, sds, os_=1.0, ls=10.0, kernel="Matern"):
    y, H = _cube(y)
    fstar, _ = _cube(fstar)
    mu_star, _ = _cube(mu_star)
    thr, _ = _cube(thr)
    n, m, _ = y.shape
    theta = _F(np.asarray(theta, dtype=np.float64).reshape(n, H))
    out = np.empty((n, H), order="F")
    _ck(lib().ref_draw_theta(C.c_uint32(seed), _u(n), _u(m), _u(H), _u(thr.shape[1] - 1), _p(y), _p(theta), _p(_F(tps)), _p(fstar), _p(mu_star),
                             _p(thr), C.c_double(os_), C.c_double(ls), kernel.encode(), _p(_F(sds)), _p(out)))
    return out


def draw_beta(seed, beta, theta, y, f, prior_means, prior_sds, thr):
    y, H = _cube(y)
    f, _ = _cube(f)
    beta, _ = _cube(beta)
    thr, _ = _cube(thr)
    n, m, _ = y.shape
    theta = _F(np.asarray(theta, dtype=np.float64).reshape(n, H))
    out = np.empty((3, m, H), order="F")
    _ck(lib().ref_draw_beta(C.c_uint32(seed), _u(n), _u(m), _u(H), _u(thr.shape[1] - 1), _p(beta), _p(theta), _p(y), _p(f), _p(_F(prior_means)),
                            _p(_F(prior_sds)), _p(thr), _p(out)))
    return out


def draw_threshold(seed, thr, y, f, mu, constant_IRF=0):
    y, H = _cube(y)
    f, _ = _cube(f)
    mu, _ = _cube(mu)
    thr, _ = _cube(thr)
    n, m, _ = y.shape
    out = np.empty_like(thr, order="F")
    _ck(lib().ref_draw_threshold(C.c_uint32(seed), _u(n), _u(m), _u(H), _u(thr.shape[1] - 1), _p(thr), _p(y), _p(f), _p(mu), C.c_int(constant_IRF),
                                 _p(out)))
    return out


def gpirtMCMC(y, theta, sample_iterations, burn_iterations, THIN, beta_prior_means, beta_prior_sds, theta_prior_means, theta_prior_sds,
              theta_os, theta_ls, KERNEL, thresholds, constant_IRF, init_normals, store_f=False, store_fstar=False):
    """The reference's gpirtMCMC() (src/gpirtMCMC.cpp:25-406). init_normals: the N(0,1) stream R::rnorm consumes during
    initialisation (3m for beta0 then n per item and session for f0)."""
    y, H = _cube(y)
    thr, _ = _cube(thresholds)
    n, m, _ = y.shape
    Cc = thr.shape[1] - 1
    theta = _F(np.asarray(theta, dtype=np.float64).reshape(n, H))
    ns = sample_iterations // THIN
    z = _F(init_normals)
    out = dict(theta=np.empty((ns, n, H), order="F"), beta=np.empty((3, m, H, ns), order="F"), threshold=np.empty((m, Cc + 1, H, ns), order="F"),
               ll=np.empty(ns), f=np.empty((n, m, H, ns), order="F") if store_f else None,
               fstar=np.empty((NGRID, m, H, ns), order="F") if store_fstar else None)
    _ck(lib().ref_gpirtMCMC(_u(n), _u(m), _u(H), _u(Cc), _p(y), _p(theta), C.c_int(sample_iterations), C.c_int(burn_iterations), C.c_int(THIN),
                            _p(_F(beta_prior_means)), _p(_F(beta_prior_sds)), _p(_F(theta_prior_means)), _p(_F(theta_prior_sds)),
                            C.c_double(theta_os), C.c_double(theta_ls), KERNEL.encode(), _p(thr), C.c_int(constant_IRF), _p(z), _u(z.size),
                            _p(out["theta"]), _p(out["beta"]), _p(out["threshold"]), _p(out["ll"]), _p(out["f"]), _p(out["fstar"])))
    return out


def recover_fstar(seed, f, y, theta, beta2, thresholds, beta_prior_means, beta_prior_sds, constant_IRF=0):
    f, H = _cube(f)
    y, _ = _cube(y)
    beta2, _ = _cube(beta2)
    thr, _ = _cube(thresholds)
    n, m, _ = f.shape
    theta = _F(np.asarray(theta, dtype=np.float64).reshape(n, H))
    bps = _F(beta_prior_sds)
    out = np.empty((NGRID, m, H), order="F")
    _ck(lib().ref_recover_fstar(C.c_int(seed), _u(n), _u(m), _u(H), _u(thr.shape[1] - 1), _p(f), _p(y), _p(theta), _p(beta2), _p(thr),
                                _p(_F(beta_prior_means)), _p(bps), _u(bps.shape[0]), C.c_int(constant_IRF), _p(out)))
    return out
