"""Host-side mirror of the reference's interface for the sweep path, over the C-ABI.

`Sampler` exposes the per-block subsystems with the reference's names (draw_f, draw_fstar, draw_theta,
draw_beta, draw_threshold, update_cholesky_cache, ll_bar — prototypes src/gpirt.h:215-290 of the
reference); `gpirtMCMC` and `recover_fstar` mirror the two R entry points (R/gpirtMCMC.R:98-167,
R/recover_fstar.R:36-46). Arrays follow the reference's shapes: y is (n, m) with NaN missing,
beta (3, m), thresholds (m, C+1), f (n, m), fstar (1001, m). No arithmetic happens here.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import NGRID, Config, GpirtError, McmcArgs, RecoverArgs, Streams, check


def _F(a, dtype=np.float64):
    return None if a is None else np.asfortranarray(np.asarray(a, dtype=dtype))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _streams(z=None, u=None, nu=None):
    """z: (nz, units), u: (ustride, units), nu: (n, m) — column-major host arrays (parity mode)."""
    if z is None and u is None and nu is None:
        return None, ()
    s = Streams()
    keep = [_F(z), _F(u), _F(nu)]
    if keep[0] is not None:
        s.z, s.z_stride = keep[0].ctypes.data, keep[0].shape[0]
    if keep[1] is not None:
        s.u, s.u_stride = keep[1].ctypes.data, keep[1].shape[0]
    if keep[2] is not None:
        s.nu = keep[2].ctypes.data
    return s, keep


def qnorm(p):
    """Standard normal quantile for the default thresholds qnorm(i/C) of R/gpirtMCMC.R:151 (Acklam + one Halley step)."""
    p = np.asarray(p, dtype=np.float64)
    from math import erfc, exp, pi, sqrt

    def one(pp):
        a = [-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02, 1.383577518672690e+02, -3.066479806614716e+01, 2.506628277459239e+00]
        b = [-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02, 6.680131188771972e+01, -1.328068155288572e+01]
        c = [-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00, -2.549732539343734e+00, 4.374664141464968e+00, 2.938163982698783e+00]
        d = [7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00, 3.754408661907416e+00]
        if pp < 0.02425:
            q = sqrt(-2 * np.log(pp))
            x = (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1)
        elif pp > 1 - 0.02425:
            q = sqrt(-2 * np.log(1 - pp))
            x = -(((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1)
        else:
            q = pp - 0.5
            r = q * q
            x = (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q / (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1)
        for _ in range(2):
            e = 0.5 * erfc(-x / sqrt(2)) - pp
            u = e * sqrt(2 * pi) * exp(x * x / 2)
            x = x - u / (1 + x * u / 2)
        return x

    return np.vectorize(one)(p)


def default_thresholds(m, Ccat):
    """thresholds[j,] = (-Inf, qnorm(1/C), ..., qnorm((C-1)/C), Inf)  (R/gpirtMCMC.R:146-154)"""
    t = np.empty((m, Ccat + 1), order="F")
    t[:, 0] = -np.inf
    t[:, Ccat] = np.inf
    for i in range(1, Ccat):
        t[:, i] = qnorm(i / Ccat)
    return t


class Sampler:
    """One chain's (item shard of the) sampler state, resident on one B200."""

    def __init__(self, n, m, Ccat, *, p_beta=3, theta_os=1.0, theta_ls=10.0, KERNEL="Matern", horizon=1, constant_IRF=0,
                 seed=0, chain=0, device=0, rank=0, world=1):
        self.n, self.m, self.C = int(n), int(m), int(Ccat)
        self.p_beta, self.rank, self.world = p_beta, rank, world
        if KERNEL not in ("Matern", "RBF"):
            raise GpirtError(_lib.ERR_INVALID, f"KERNEL {KERNEL!r}: the reference only knows 'Matern' and 'RBF'")
        cfg = Config(n=self.n, m_total=self.m, horizon=horizon, C=self.C, constant_IRF=constant_IRF, p_beta=p_beta,
                     theta_os=theta_os, theta_ls=theta_ls, kernel=0 if KERNEL == "Matern" else 1, device=device,
                     seed=seed, chain=chain, rank=rank, world=world)
        self._L = _lib.lib()
        h = C.c_void_p()
        rc = self._L.gpirt_create(C.byref(cfg), C.byref(h))
        self._h = h if h.value else None
        if rc != _lib.OK:
            msg = self._L.gpirt_last_error().decode()
            self.close()
            raise GpirtError(rc, msg)
        self.m_chunk = (self.m + world - 1) // world
        self.n_chunk = (self.n + world - 1) // world

    @classmethod
    def _borrowed(cls, handle, n, m, Ccat, p_beta=3):
        """A Sampler over a handle somebody else owns (a session of a Panel): close() forgets it instead of destroying it."""
        self = cls.__new__(cls)
        self.n, self.m, self.C, self.p_beta, self.rank, self.world = int(n), int(m), int(Ccat), p_beta, 0, 1
        self._L, self._h, self._owned = _lib.lib(), handle, False
        self.m_chunk, self.n_chunk = self.m, self.n
        return self

    def close(self):
        if getattr(self, "_h", None):
            if getattr(self, "_owned", True):
                self._L.gpirt_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- data / state ----
    def set_data(self, y):
        y = np.asarray(y)
        if y.shape != (self.n, self.m):
            raise GpirtError(_lib.ERR_SHAPE, f"y has shape {y.shape}, expected {(self.n, self.m)}")
        if y.dtype == np.int8:
            yy = _F(y, np.int8)
            check(self._L.gpirt_set_data_i8(self._h, _p(yy)))
        else:
            yy = _F(y)
            check(self._L.gpirt_set_data_f64(self._h, _p(yy)))

    def set_mean_rows(self, mu):
        """per-row means for draw_f / draw_threshold (the stacked sampler of a constant-IRF panel); None returns to X beta"""
        if mu is None:
            check(self._L.gpirt_set_mean_rows(self._h, None))
            return
        mu = _F(mu)
        if mu.shape != (self.n, self.m):
            raise GpirtError(_lib.ERR_SHAPE, f"mu has shape {mu.shape}, expected {(self.n, self.m)}")
        check(self._L.gpirt_set_mean_rows(self._h, _p(mu)))

    def set_priors(self, beta_prior_sds=None, theta_prior_sds=None):
        b, t = _F(beta_prior_sds), _F(theta_prior_sds)
        if b is not None and b.shape != (3, self.m):
            raise GpirtError(_lib.ERR_SHAPE, f"beta_prior_sds has shape {b.shape}, expected {(3, self.m)}")
        if t is not None and t.shape != (2, self.n):
            raise GpirtError(_lib.ERR_SHAPE, f"theta_prior_sds has shape {t.shape}, expected {(2, self.n)}")
        check(self._L.gpirt_set_priors(self._h, _p(b), _p(t)))

    def set_state(self, theta=None, beta=None, f=None, thresholds=None, fstar=None):
        shapes = dict(theta=(self.n,), beta=(self.p_beta, self.m), f=(self.n, self.m), thresholds=(self.m, self.C + 1), fstar=(NGRID, self.m))
        arrs = dict(theta=_F(theta), beta=_F(beta), f=_F(f), thresholds=_F(thresholds), fstar=_F(fstar))
        for k, a in arrs.items():
            if a is not None and a.size != int(np.prod(shapes[k])):
                raise GpirtError(_lib.ERR_SHAPE, f"{k} has shape {a.shape}, expected {shapes[k]}")
        check(self._L.gpirt_set_state(self._h, _p(arrs["theta"]), _p(arrs["beta"]), _p(arrs["f"]), _p(arrs["thresholds"]), _p(arrs["fstar"])))

    def get_state(self, theta=True, beta=True, f=True, thresholds=True, fstar=False):
        out = {}
        if theta:
            out["theta"] = np.empty(self.n)
        if beta:
            out["beta"] = np.empty((self.p_beta, self.m), order="F")
        if f:
            out["f"] = np.zeros((self.n, self.m), order="F")
        if thresholds:
            out["thresholds"] = np.empty((self.m, self.C + 1), order="F")
        if fstar:
            out["fstar"] = np.empty((NGRID, self.m), order="F")
        check(self._L.gpirt_get_state(self._h, _p(out.get("theta")), _p(out.get("beta")), _p(out.get("f")), _p(out.get("thresholds")),
                                      _p(out.get("fstar"))))
        return out

    def get_L(self):
        out = np.empty((self.n, self.n), order="F")
        check(self._L.gpirt_get_L(self._h, _p(out)))
        return out

    # ---- device-side store of kept draws (src/gpirtMCMC.cpp:230-255,334-356) ----
    def store_open(self, capacity, store_f=False, store_fstar=False):
        check(self._L.gpirt_store_open(self._h, int(capacity), int(store_f), int(store_fstar)))
        self._store_flags = (bool(store_f), bool(store_fstar))

    def store_push(self):
        """Appends the current state as the next sample; no host synchronisation."""
        check(self._L.gpirt_store_push(self._h))

    def store_count(self):
        cnt, cap = C.c_int(), C.c_int()
        check(self._L.gpirt_store_count(self._h, C.byref(cnt), C.byref(cap)))
        return cnt.value, cap.value

    def store_sample_bytes(self, store_f=False, store_fstar=False):
        return int(self._L.gpirt_store_sample_bytes(self._h, int(store_f), int(store_fstar)))

    def store_flush(self, out, first):
        """Writes the buffered samples as samples first.. of the reference-layout arrays in `out` (keys theta [n_samples, n, 1], beta
        [3, m, 1, n_samples], threshold [m, C+1, 1, n_samples], ll [n_samples], optional f, fstar), all Fortran-ordered."""
        n_samples = out["ll"].shape[0]
        check(self._L.gpirt_store_flush(self._h, int(n_samples), int(first), _p(out.get("theta")), _p(out.get("beta")), _p(out.get("threshold")),
                                        _p(out.get("ll")), _p(out.get("f")), _p(out.get("fstar"))))

    def store_close(self):
        check(self._L.gpirt_store_close(self._h))

    def set_iter(self, it):
        check(self._L.gpirt_set_iter(self._h, C.c_uint32(it)))

    def set_solver(self, solver):
        """0: blocked substitution for L^-1 k* / L^-T(.) in draw_fstar; 1: explicit block inverse + GEMMs (default)"""
        check(self._L.gpirt_set_solver(self._h, int(solver)))

    def set_overlap(self, on):
        """run the theta-only chain of the next sweep on a second stream under the item samplers (default on)"""
        check(self._L.gpirt_set_overlap(self._h, int(on)))

    def set_screen(self, on):
        """fp32 screening of ESS proposals with a rigorous error bound (default on); decisions are identical to fp64"""
        check(self._L.gpirt_set_screen(self._h, int(on)))

    def exact_evals(self, block):
        """per unit: evaluations of the last call of `block` that ran in fp64"""
        units = self.n if block == "theta" else self.m
        out = np.zeros(units, dtype=np.int32)
        check(self._L.gpirt_get_exact_evals(self._h, _lib.BLOCK[block], _p(out)))
        return out

    def set_stream(self, cuda_stream):
        check(self._L.gpirt_set_stream(self._h, C.c_void_p(cuda_stream)))

    def sync(self):
        check(self._L.gpirt_sync(self._h))

    def diag(self, block):
        units = self.n if block == "theta" else self.m
        ev = np.empty(units, dtype=np.int32)
        mg = np.empty(units)
        check(self._L.gpirt_get_diag(self._h, _lib.BLOCK[block], _p(ev), _p(mg)))
        return ev, mg

    def init_draw(self, beta_prior_means=None, draw_beta=True, draw_f=True):
        b = _F(beta_prior_means)
        check(self._L.gpirt_init_draw(self._h, _p(b), int(draw_beta), int(draw_f)))

    def init_f_from_z(self, z):
        """f_j = L z_j for every local item (rmvnorm(L), src/mvnormal.h:7-14; the reference's initial f, src/gpirtMCMC.cpp:158-162)."""
        zz = _F(z)
        if zz.shape != (self.n, self.m):
            raise GpirtError(_lib.ERR_SHAPE, f"z has shape {zz.shape}, expected {(self.n, self.m)}")
        check(self._L.gpirt_init_f_from_z(self._h, _p(zz)))

    # ---- subsystems (reference names) ----
    def update_cholesky_cache(self):
        check(self._L.gpirt_update_cholesky_cache(self._h))

    def _block(self, fn, z=None, u=None, nu=None):
        s, keep = _streams(z, u, nu)
        check(fn(self._h, C.byref(s) if s is not None else None))
        if keep:
            self.sync()  # the staged host arrays must outlive the copy

    def draw_f(self, z=None, u=None, nu=None):
        self._block(self._L.gpirt_draw_f, z, u, nu)

    def draw_fstar(self, z=None):
        self._block(self._L.gpirt_draw_fstar, z)

    def draw_theta(self, z=None, u=None):
        self._block(self._L.gpirt_draw_theta, z, u)

    def gather_f(self):
        check(self._L.gpirt_gather_f(self._h))

    def draw_beta(self, z=None, u=None):
        self._block(self._L.gpirt_draw_beta, z, u)

    def draw_threshold(self, z=None, u=None):
        self._block(self._L.gpirt_draw_threshold, z, u)

    def ll_bar(self):
        per = np.empty(self.m)
        tot = C.c_double()
        check(self._L.gpirt_ll_bar(self._h, _p(per), C.byref(tot)))
        return per, tot.value

    def total_ll(self):
        tot = C.c_double()
        check(self._L.gpirt_total_ll(self._h, C.byref(tot)))
        return tot.value

    def sweep(self, k=1):
        check(self._L.gpirt_sweep(self._h, int(k)))

    def phase_items(self):
        check(self._L.gpirt_sweep_phase_items(self._h))

    def phase_theta(self):
        check(self._L.gpirt_sweep_phase_theta(self._h))

    def phase_close(self):
        check(self._L.gpirt_sweep_phase_close(self._h))

    # ---- item-sharded theta step (include/gpirt_b200.h): begin, then rounds of partials / exchange / decide ----
    def theta_items_begin(self, z=None, u=None):
        s, keep = _streams(z, u)
        check(self._L.gpirt_theta_items_begin(self._h, C.byref(s) if s is not None else None))
        if keep:
            self.sync()

    def theta_items_partials(self):
        check(self._L.gpirt_theta_items_partials(self._h))

    def theta_items_decide(self):
        n_active = C.c_int32()
        check(self._L.gpirt_theta_items_decide(self._h, C.byref(n_active)))
        return n_active.value

    def items_pack(self, unpack=False):
        check(self._L.gpirt_items_pack(self._h, int(unpack)))

    # ---- f* exchange through NVLink peer memory (item shards, one process per GPU) ----
    def ipc_export(self):
        buf, o1, o2 = (C.c_ubyte * 64)(), C.c_int64(), C.c_int64()
        check(self._L.gpirt_ipc_export(self._h, buf, C.byref(o1), C.byref(o2)))
        return bytes(buf), o1.value, o2.value

    def ipc_attach(self, peer_rank, exported):
        handle, o1, o2 = exported
        buf = (C.c_ubyte * 64).from_buffer_copy(handle)
        check(self._L.gpirt_ipc_attach(self._h, int(peer_rank), buf, C.c_int64(o1), C.c_int64(o2)))

    def ipc_detach(self):
        check(self._L.gpirt_ipc_detach(self._h))

    def peer_wait(self):
        check(self._L.gpirt_peer_wait(self._h))

    # ---- distributed theta-only chain (n >= 4096 on item shards): the peers' chain buffers ----
    def chain_export(self):
        buf, offs = (C.c_ubyte * 64)(), (C.c_int64 * 6)()
        check(self._L.gpirt_chain_export(self._h, buf, offs))
        return bytes(buf), list(offs)

    def chain_attach(self, peer_rank, exported, same_process=False):
        handle, offs = exported
        o = (C.c_int64 * 6)(*offs)
        buf = None if same_process else (C.c_ubyte * 64).from_buffer_copy(handle)
        check(self._L.gpirt_chain_attach(self._h, int(peer_rank), buf, o))

    def chain_detach(self):
        check(self._L.gpirt_chain_detach(self._h))

    def exchange_buffer(self, which):
        ptr, nbytes = C.c_void_p(), C.c_int64()
        check(self._L.gpirt_exchange_buffer(self._h, _lib.XCHG[which], C.byref(ptr), C.byref(nbytes)))
        return ptr.value, nbytes.value

    # ---- measurement / test hooks ----
    def profile(self, on=True):
        check(self._L.gpirt_profile_enable(self._h, int(on)))

    def profile_reset(self):
        check(self._L.gpirt_profile_reset(self._h))

    def profile_read(self):
        """{section: (total_ms, launches)} accumulated since the last reset"""
        cap = 64
        names = C.create_string_buffer(32 * cap)
        ms = (C.c_double * cap)()
        cnt = (C.c_int64 * cap)()
        k = C.c_int()
        check(self._L.gpirt_profile_read(self._h, names, ms, cnt, cap, C.byref(k)))
        return {names.raw[32 * i:32 * i + 32].split(b"\0")[0].decode(): (ms[i], cnt[i]) for i in range(k.value)}

    def profile_timeline(self, cap=512):
        """[(section, begin_ms, end_ms)] of everything recorded since the last read / reset, in issue order, on the device clock (profiling level 2)"""
        names = C.create_string_buffer(32 * cap)
        b, e = (C.c_double * cap)(), (C.c_double * cap)()
        k = C.c_int()
        check(self._L.gpirt_profile_timeline(self._h, names, b, e, cap, C.byref(k)))
        return [(names.raw[32 * i:32 * i + 32].split(b"\0")[0].decode(), b[i], e[i]) for i in range(k.value)]

    def debug_uniforms(self, stream, unit, count):
        out = np.empty(count)
        check(self._L.gpirt_debug_uniforms(self._h, C.c_uint32(stream), C.c_uint32(unit), int(count), _p(out)))
        return out

    def debug_normals(self, stream, unit0, rows, cols):
        out = np.empty((rows, cols), order="F")
        check(self._L.gpirt_debug_normals(self._h, C.c_uint32(stream), C.c_uint32(unit0), int(rows), int(cols), _p(out)))
        return out
