"""TEST INFRASTRUCTURE — ctypes front-end to the CPU oracle (oracle/_build/liborc.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product package gpirt_b200 never does. All matrices are column-major float64,
exactly as the reference's Armadillo containers lay them out (y: n x m with NaN = missing).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liborc.so")
NGRID = 1001
BLOCKS = {"f": 0, "fstar": 1, "theta": 2, "beta": 3, "thr": 4}


def build(force=False):
    """Compile liborc.so (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(_LIB):
        subprocess.run(["make", "-C", _HERE, "_build/liborc.so"], check=True, capture_output=True)
    return _LIB


class Src(C.Structure):
    _fields_ = [("mode", C.c_int32), ("z", C.c_void_p), ("zstride", C.c_uint64), ("u", C.c_void_p),
                ("ustride", C.c_uint64), ("seed", C.c_uint64), ("chain", C.c_uint32), ("iter", C.c_uint32),
                ("session", C.c_uint32), ("pad_", C.c_uint32)]


_lib = None
_dp = C.POINTER(C.c_double)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        L.orc_pnorm.restype = C.c_double
        L.orc_pnorm.argtypes = [C.c_double]
        L.orc_ll_bar_sparse.restype = C.c_double
        L.orc_grid_value.restype = C.c_double
        L.orc_grid_value.argtypes = [C.c_int]
        L.orc_grid_index.argtypes = [C.c_double]
        L.orc_total_ll.restype = C.c_double
        L.orc_last_ll.restype = C.c_double
        L.orc_create.restype = C.c_void_p
        for name in ("orc_destroy", "orc_set_state", "orc_get_state", "orc_get_L", "orc_set_L", "orc_get_mu",
                     "orc_compute_mu", "orc_update_cholesky", "orc_mt_seed_all", "orc_num_threads", "orc_draw_f",
                     "orc_draw_fstar", "orc_draw_theta", "orc_gather_f", "orc_draw_beta", "orc_draw_threshold",
                     "orc_total_ll", "orc_item_ll", "orc_sweep", "orc_last_ll", "orc_get_diag", "orc_run_mt",
                     "orc_set_opts"):
            getattr(L, name).argtypes = None
        _lib = L
    return _lib


def find_openblas():
    """The OpenBLAS inside the scipy wheel of this image (LP64, symbols prefixed scipy_), or None."""
    import glob
    import importlib.util

    spec = importlib.util.find_spec("scipy")
    if spec is None or not spec.submodule_search_locations:
        return None
    base = os.path.dirname(list(spec.submodule_search_locations)[0])
    hits = sorted(glob.glob(os.path.join(base, "scipy.libs", "libscipy_openblas-*.so")))
    return hits[0] if hits else None


def use_blas(on=True, threads=None):
    """TIMED BASELINE ONLY (oracle/orc_blas.h): route chol / triangular solves / gemv of the oracle through an optimised BLAS, as the
    reference's Armadillo calls would be in an R linked against one. Returns the library path, or None when it is unavailable."""
    if not on:
        lib().orc_blas_off()
        return None
    path = find_openblas()
    if path is None:
        return None
    rc = lib().orc_blas_load(path.encode(), C.c_int(int(threads or os.cpu_count() or 1)))
    return path if rc == 0 else None


def set_threads(k):
    """OpenMP threads of samplers created after this call; 1 reproduces the reference's single-thread RNG order."""
    lib().orc_set_threads(C.c_int(int(k)))


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _F(a):
    """column-major float64 copy"""
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


# ---- leaves ------------------------------------------------------------------------------------
def pnorm(x):
    x = _f64(x).ravel()
    out = np.empty_like(x)
    lib().orc_pnorm_vec(_p(x), C.c_uint64(x.size), _p(out))
    return out


def K(x1, x2, sds):
    x1, x2, sds = _f64(x1), _f64(x2), _f64(sds)
    out = np.empty((x1.size, x2.size), order="F")
    lib().orc_K(_p(x1), C.c_uint64(x1.size), _p(x2), C.c_uint64(x2.size), _p(sds), _p(out))
    return out


def K_time(x1, x2, os_, ls, tsds, kernel="Matern"):
    x1, x2, tsds = _f64(x1), _f64(x2), _f64(tsds)
    out = np.zeros((x1.size, x2.size), order="F")
    k = {"Matern": 0, "RBF": 1}.get(kernel, 2)
    lib().orc_K_time(_p(x1), C.c_uint64(x1.size), _p(x2), C.c_uint64(x2.size), C.c_double(os_), C.c_double(ls),
                     _p(tsds), C.c_int(k), _p(out))
    return out


def chol(A, long_double=False):
    A = _F(A).copy(order="F")
    fn = lib().orc_chol_ld if long_double else lib().orc_chol
    info = fn(_p(A), C.c_uint64(A.shape[0]))
    return A, info


def double_solve(L, x):
    L = _F(L)
    x = _f64(x).copy()
    lib().orc_double_solve(_p(L), C.c_uint64(L.shape[0]), _p(x))
    return x


def trsv_lower(L, x):
    L = _F(L)
    x = _f64(x).copy()
    lib().orc_trsv_lower(_p(L), C.c_uint64(L.shape[0]), _p(x))
    return x


def ll_bar_sparse(f, y, mu, thr):
    f, y, mu, thr = _f64(f), _f64(y), _f64(mu), _f64(thr)
    return lib().orc_ll_bar_sparse(_p(f), _p(y), _p(mu), _p(thr), C.c_uint64(f.size))


def delta_to_threshold(d):
    d = _f64(d)
    Cc = d.size + 1
    out = np.empty(Cc + 1)
    lib().orc_delta_to_threshold(_p(d), C.c_uint64(Cc), _p(out))
    return out


def threshold_to_delta(t):
    t = _f64(t)
    Cc = t.size - 1
    out = np.empty(Cc - 1)
    lib().orc_threshold_to_delta(_p(t), C.c_uint64(Cc), _p(out))
    return out


def grid_index(theta):
    return lib().orc_grid_index(C.c_double(theta))


def grid_value(idx):
    return lib().orc_grid_value(C.c_int(idx))


def philox_uniform(seed, chain, it, stream, unit, count):
    out = np.empty(count)
    lib().orc_philox_uniform(C.c_uint64(seed), C.c_uint32(chain), C.c_uint32(it), C.c_uint32(stream),
                             C.c_uint32(unit), C.c_uint64(count), _p(out))
    return out


def philox_normal(seed, chain, it, stream, unit, count):
    out = np.empty(count)
    lib().orc_philox_normal(C.c_uint64(seed), C.c_uint32(chain), C.c_uint32(it), C.c_uint32(stream),
                            C.c_uint32(unit), C.c_uint64(count), _p(out))
    return out


def philox_raw(ctr, key):
    out = np.empty(4, dtype=np.uint32)
    lib().orc_philox_raw(*(C.c_uint32(int(c)) for c in ctr), C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), _p(out))
    return out


# ---- sampler -----------------------------------------------------------------------------------
def supplied(z=None, u=None):
    """Src for caller-supplied streams: z is (nz, units) column-major, u is (ustride, units)."""
    s = Src()
    s.mode = 0
    s._keep = []
    if z is not None:
        z = _F(z)
        s.z, s.zstride = z.ctypes.data, z.shape[0]
        s._keep.append(z)
    if u is not None:
        u = _F(u)
        s.u, s.ustride = u.ctypes.data, u.shape[0]
        s._keep.append(u)
    return s


def philox(seed, chain, it, session=0):
    s = Src()
    s.mode, s.seed, s.chain, s.iter, s.session = 1, seed, chain, it, session
    return s


def mt():
    s = Src()
    s.mode = 2
    return s


class Sampler:
    """CPU oracle of the reference sweep at horizon 1, constant_IRF 0 (src/gpirtMCMC.cpp:261-331)."""

    def __init__(self, y, C_, beta_prior_sds, theta_prior_sds=None, faithful=0, p_beta=3, theta_regime=0, theta_os=1.0):
        y = _F(y)
        self.n, self.m = y.shape
        self.C = int(C_)
        sds = _F(beta_prior_sds)
        assert sds.shape == (3, self.m)
        tps = _F(np.zeros((2, self.n)) if theta_prior_sds is None else theta_prior_sds)
        self._h = C.c_void_p(lib().orc_create(C.c_uint64(self.n), C.c_uint64(self.m), C.c_uint64(self.C), _p(y), _p(sds), _p(tps)))
        lib().orc_set_opts(self._h, C.c_int(faithful), C.c_int(p_beta), C.c_int(theta_regime), C.c_double(theta_os))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_destroy(self._h)
            self._h = None

    def set_state(self, theta=None, beta=None, f=None, thr=None, fstar=None):
        a = [None if x is None else _F(x) for x in (theta, beta, f, thr, fstar)]
        if a[1] is not None and a[1].shape[0] == 2:  # recover_fstar's 2-row beta
            a[1] = _F(np.vstack([a[1], np.zeros((1, self.m))]))
        lib().orc_set_state(self._h, *[_p(x) for x in a])

    def state(self):
        th = np.empty(self.n)
        be = np.empty((3, self.m), order="F")
        f = np.empty((self.n, self.m), order="F")
        thr = np.empty((self.m, self.C + 1), order="F")
        fs = np.empty((NGRID, self.m), order="F")
        lib().orc_get_state(self._h, _p(th), _p(be), _p(f), _p(thr), _p(fs))
        return dict(theta=th, beta=be, f=f, thr=thr, fstar=fs)

    def L(self):
        out = np.empty((self.n, self.n), order="F")
        lib().orc_get_L(self._h, _p(out))
        return out

    def set_L(self, L):
        L = _F(L)
        lib().orc_set_L(self._h, _p(L))

    def mu(self):
        mu = np.empty((self.n, self.m), order="F")
        ms = np.empty((NGRID, self.m), order="F")
        lib().orc_get_mu(self._h, _p(mu), _p(ms))
        return mu, ms

    def compute_mu(self):
        lib().orc_compute_mu(self._h)

    def update_cholesky(self):
        return lib().orc_update_cholesky(self._h)

    def mt_seed_all(self, base):
        lib().orc_mt_seed_all(self._h, C.c_uint32(base))

    def num_threads(self):
        return lib().orc_num_threads(self._h)

    def draw_f(self, src, nu=None):
        nu = None if nu is None else _F(nu)
        lib().orc_draw_f(self._h, C.byref(src), _p(nu))

    def draw_fstar(self, src):
        return lib().orc_draw_fstar(self._h, C.byref(src))

    def draw_theta(self, src):
        lib().orc_draw_theta(self._h, C.byref(src))

    def gather_f(self):
        lib().orc_gather_f(self._h)

    def draw_beta(self, src):
        lib().orc_draw_beta(self._h, C.byref(src))

    def draw_threshold(self, src):
        lib().orc_draw_threshold(self._h, C.byref(src))

    def total_ll(self):
        return lib().orc_total_ll(self._h)

    def item_ll(self):
        out = np.empty(self.m)
        lib().orc_item_ll(self._h, _p(out))
        return out

    def sweep(self, srcs):
        arr = (Src * 5)()
        keep = []
        for k, name in enumerate(("f", "fstar", "theta", "beta", "thr")):
            s = srcs[name] if isinstance(srcs, dict) else srcs
            for fld, _ in Src._fields_:
                setattr(arr[k], fld, getattr(s, fld))
            keep.append(s)
        return lib().orc_sweep(self._h, arr)

    def times(self, reset=True):
        """(seconds in work shared by all items, seconds in per-item / per-respondent work) since the last reset"""
        a, b = C.c_double(), C.c_double()
        lib().orc_get_times(self._h, C.byref(a), C.byref(b), C.c_int(int(reset)))
        return a.value, b.value

    def last_ll(self):
        return lib().orc_last_ll(self._h)

    def diag(self, block):
        units = self.n if block == "theta" else self.m
        ev = np.empty(units, dtype=np.int32)
        mg = np.empty(units)
        over = lib().orc_get_diag(self._h, C.c_int(BLOCKS[block]), _p(ev), _p(mg))
        return ev, mg, bool(over)

    def record(self, on=True):
        """keep every unit's normals / uniforms of the next block calls (see recorded())"""
        lib().orc_record(self._h, C.c_int(int(on)))

    def recorded(self, block, zcap, ucap=64):
        """(z zcap x units, u ucap x units, nz, nu) drawn by the last call of `block`, zero padded"""
        units = self.n if block == "theta" else self.m
        z = np.zeros((zcap, units), order="F")
        u = np.zeros((ucap, units), order="F")
        nz = np.zeros(units, dtype=np.int32)
        nu = np.zeros(units, dtype=np.int32)
        over = lib().orc_get_record(self._h, C.c_int(BLOCKS[block]), _p(z), C.c_uint64(zcap), _p(u), C.c_uint64(ucap), _p(nz), _p(nu))
        assert not over, "recorded stream longer than the requested capacity"
        return z, u, nz, nu

    def run_mt(self, iter0, iters, store=True):
        th = np.empty((self.n, iters), order="F") if store else None
        ll = np.empty(iters) if store else None
        info = lib().orc_run_mt(self._h, C.c_int(iter0), C.c_int(iters), _p(th), _p(ll))
        return info, th, ll


class Panel:
    """CPU oracle at horizon > 1, constant_IRF 0: H single-session Samplers (sessions are independent given theta) plus the
    reference's joint theta step (src/draw-theta.cpp:171-236): RDM (theta_ls <= 0.1) or GP over time (0.1 < theta_ls < 3H).
    The CST regime (theta_ls >= 3H) is an out-of-bounds read of the thresholds in the reference at H > 1 and is refused.
    Arrays follow the reference's cubes: y n x m x H, theta n x H, beta 3 x m x H, f n x m x H, thr m x (C+1) x H, fstar NGRID x m x H."""

    def __init__(self, y, C_, beta_prior_sds, theta_prior_sds=None, theta_os=1.0, theta_ls=10.0, KERNEL="Matern"):
        y = np.asfortranarray(np.asarray(y, dtype=np.float64))
        self.n, self.m, self.H = y.shape
        self.C = int(C_)
        H = self.H
        if theta_ls >= 3 * H and H > 1:
            raise ValueError("CST regime at horizon > 1: the reference reads thresholds out of bounds (src/draw-theta.cpp:162)")
        # regime 0: one session in the CST regime (theta_ls >= 3, R's default 10): the session's own theta step (src/draw-theta.cpp:112-170)
        self.regime = 0 if theta_ls >= 3 * H else (1 if theta_ls <= 0.1 else 2)
        self.S = [Sampler(y[:, :, h], C_, beta_prior_sds, theta_prior_sds) for h in range(H)]
        for h in range(1, H):
            lib().orc_share_mts(self.S[h]._h, self.S[0]._h)
        ts = np.arange(float(H))  # linspace(0, H-1, H), src/cholesky-cache.cpp:28
        V = K_time(ts, ts, theta_os, theta_ls, [0.0, 0.0], KERNEL) + 1e-6 * np.eye(H)  # :29-32 (theta_prior_sds is NOT used there)
        self.L_time, info = chol(V)
        assert info == 0
        self.L_time = _F(self.L_time)

    def set_state(self, theta=None, beta=None, f=None, thr=None, fstar=None):
        for h, S in enumerate(self.S):
            S.set_state(theta=None if theta is None else np.asarray(theta)[:, h], beta=None if beta is None else np.asarray(beta)[:, :, h],
                        f=None if f is None else np.asarray(f)[:, :, h], thr=None if thr is None else np.asarray(thr)[:, :, h],
                        fstar=None if fstar is None else np.asarray(fstar)[:, :, h])

    def state(self):
        st = [S.state() for S in self.S]
        return dict(theta=np.stack([s["theta"] for s in st], axis=1), beta=np.stack([s["beta"] for s in st], axis=2),
                    f=np.stack([s["f"] for s in st], axis=2), thr=np.stack([s["thr"] for s in st], axis=2),
                    fstar=np.stack([s["fstar"] for s in st], axis=2))

    def mt_seed_all(self, base):
        self.S[0].mt_seed_all(base)  # the one pool every session consumes

    def record(self, on=True):
        for S in self.S:
            S.record(on)

    def compute_mu(self):
        for S in self.S:
            S.compute_mu()

    def update_cholesky(self):
        return max(S.update_cholesky() for S in self.S)

    def draw_theta(self, srcs):
        """srcs: one Src per session (RDM) or a single Src (GP; H normals per respondent)"""
        srcs = list(srcs) if isinstance(srcs, (list, tuple)) else [srcs]
        if self.regime == 0:
            self.S[0].draw_theta(srcs[0])
            return
        arr = (Src * len(srcs))()
        for k, s in enumerate(srcs):
            for fld, _ in Src._fields_:
                setattr(arr[k], fld, getattr(s, fld))
        hs = (C.c_void_p * self.H)(*[S._h for S in self.S])
        lib().orc_panel_draw_theta(hs, C.c_uint64(self.H), C.c_int(self.regime), _p(self.L_time), arr)
        self._keep = srcs

    def sweep(self, src_of):
        """One sweep in the reference's order (src/gpirtMCMC.cpp:267-331). src_of(block, h) -> Src for session h (h is None for the GP theta step)."""
        for S in self.S:  # draw_f calls update_cholesky_cache itself (src/draw-f.cpp:59)
            S.compute_mu()
            if S.update_cholesky() != 0:
                return 1
        for h, S in enumerate(self.S):
            S.draw_f(src_of("f", h))
        for h, S in enumerate(self.S):
            if S.draw_fstar(src_of("fstar", h)) != 0:
                return 2
        self.draw_theta([src_of("theta", h) for h in range(self.H)] if self.regime != 2 else src_of("theta", None))
        for S in self.S:
            S.gather_f()
        for h, S in enumerate(self.S):
            S.draw_beta(src_of("beta", h))
        for S in self.S:
            S.compute_mu()
            if S.update_cholesky() != 0:
                return 1
        for h, S in enumerate(self.S):
            S.draw_threshold(src_of("thr", h))
        return 0

    def total_ll(self):
        ll = 0.0
        for S in self.S:  # src/gpirtMCMC.cpp:324-331: one running sum over (h, j); per session here, then across sessions in order
            ll += S.total_ll()
        return ll


class ConstantPanel(Panel):
    """CPU oracle at constant_IRF == 1 (one item response function shared by all sessions), restated from src/draw-f.cpp:84-139,
    src/draw-fstar.cpp:58-125, src/draw_threshold.cpp:5-96,181-205 and the constant branch of the initialisation (src/gpirtMCMC.cpp:164-200).
    It is the horizon > 1 panel plus one STACKED sampler T over the n*H (respondent, session) rows:
      draw_f          one ESS per item on the stacked column, nu = chol(K(theta_stacked) + 1e-6 I) z, thresholds of session 0,
                      mean = the sessions' own mu_h stacked (mu_h = X_h beta_h; in the first sweep X_h beta_0, as the initialisation leaves it)
      draw_fstar      inducing-point posterior (Sampler.draw_fstar_cst) from the stacked f, mu* of session 0, the same f* for every session
      draw_theta, f <- f* gather, draw_beta, mu / mu* update: per session, unchanged (they take no constant_IRF argument)
      draw_threshold  one ESS per item over all sessions' cells against session 0's thresholds, result copied to every session
    TEST INFRASTRUCTURE (CPU); the CUDA counterpart is gpirt_b200.ConstantPanel / mcmc_constant (csrc/entry.cu)."""

    def __init__(self, y, C_, beta_prior_sds, theta_prior_sds=None, theta_os=1.0, theta_ls=10.0, KERNEL="Matern"):
        super().__init__(y, C_, beta_prior_sds, theta_prior_sds, theta_os, theta_ls, KERNEL)
        y = np.asfortranarray(np.asarray(y, dtype=np.float64))
        n, m, H = y.shape
        ystack = np.asfortranarray(np.concatenate([y[:, :, h] for h in range(H)], axis=0))  # row h*n + i, src/draw-f.cpp:94-103
        tps = np.zeros((2, n * H)) if theta_prior_sds is None else np.tile(np.asarray(theta_prior_sds), (1, H))
        self.T = Sampler(ystack, C_, beta_prior_sds, tps)
        lib().orc_share_mts(self.T._h, self.S[0]._h)
        lib().orc_draw_fstar_cst.argtypes = None
        lib().orc_set_mu.argtypes = None

    # ---- helpers ----
    def _stack(self, key):
        return np.asfortranarray(np.concatenate([S.state()[key] if key != "mu" else S.mu()[0] for S in self.S], axis=0))

    def _load_stacked(self, thr0):
        st0 = self.S[0].state()
        theta = np.concatenate([S.state()["theta"] for S in self.S])
        self.T.set_state(theta=theta, beta=st0["beta"], f=self._stack("f"), thr=thr0)
        self.T.compute_mu()  # mu* = X* beta_0 (src/draw-fstar.cpp:114 uses mu_star.slice(0)); mu is replaced by the sessions' own below
        mu = self._stack("mu")
        lib().orc_set_mu(self.T._h, _p(mu))
        self._keep_mu = mu

    def set_mu(self, h, mu):
        mu = _F(mu)
        lib().orc_set_mu(self.S[h]._h, _p(mu))

    def init_constant(self, theta, beta0, f0, thr):
        """State after the constant branch of the initialisation: beta drawn for session 0 only (the other slices stay zero), every session's
        mu = X_h beta_0, the same f0 columns in every session (src/gpirtMCMC.cpp:164-200); mu* per session from its own beta slice (:214-216)."""
        H = self.H
        theta = np.asarray(theta)
        for h, S in enumerate(self.S):
            S.set_state(theta=theta[:, h], beta=beta0 if h == 0 else np.zeros_like(beta0), f=f0, thr=np.asarray(thr)[:, :, h])
            S.compute_mu()  # mu* = X* beta_h
            X = np.stack([np.ones(self.n), theta[:, h], theta[:, h] ** 2], axis=1)
            self.set_mu(h, X @ beta0)
        return H

    def draw_f(self, src):
        self._load_stacked(self.S[0].state()["thr"])
        assert self.T.update_cholesky() == 0  # chol(K(theta_constant) + 1e-6 I), src/draw-f.cpp:106-108
        self.T.draw_f(src)
        f = self.T.state()["f"]
        for h, S in enumerate(self.S):
            S.set_state(f=f[h * self.n:(h + 1) * self.n])

    def draw_fstar(self, src):
        self._load_stacked(self.S[0].state()["thr"])
        info = lib().orc_draw_fstar_cst(self.T._h, C.byref(src))
        fs = self.T.state()["fstar"]
        for S in self.S:
            S.set_state(fstar=fs)
        return info

    def draw_threshold(self, src):
        self._load_stacked(self.S[0].state()["thr"])
        self.T.draw_threshold(src)
        thr = self.T.state()["thr"]
        for S in self.S:
            S.set_state(thr=thr)

    def sweep(self, src_of):
        """One sweep (src/gpirtMCMC.cpp:267-331) at constant_IRF == 1. src_of(block, h): h is None for the stacked blocks and the GP theta step.
        mu is NOT recomputed at the top: it is what the previous sweep (or the initialisation) left."""
        self.draw_f(src_of("f", None))
        if self.draw_fstar(src_of("fstar", None)) != 0:
            return 2
        self.draw_theta([src_of("theta", h) for h in range(self.H)] if self.regime != 2 else src_of("theta", None))
        for S in self.S:
            S.gather_f()
        for h, S in enumerate(self.S):
            S.draw_beta(src_of("beta", h))
        for S in self.S:
            S.compute_mu()
            if S.update_cholesky() != 0:
                return 1
        self.draw_threshold(src_of("thr", None))
        return 0
