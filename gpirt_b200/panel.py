"""horizon > 1 (dynamic GPIRT), constant_IRF == 0: the reference's cubes (y n x m x H, theta n x H, beta 3 x m x H, f n x m x H,
thresholds m x (C+1) x H, fstar NGRID x m x H) over a gpirt_panel — H single-session samplers plus the joint theta step of
src/draw-theta.cpp:171-236 (RDM for theta_ls <= 0.1, GP over time for 0.1 < theta_ls < 3H; theta_ls >= 3H is refused, see
include/gpirt_b200.h)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import NGRID, Config, GpirtError, Streams, check
from .sampler import Sampler, _F, _p, _streams


class Panel:
    def __init__(self, n, m, Ccat, horizon, *, theta_os=1.0, theta_ls=10.0, KERNEL="Matern", seed=0, chain=0, device=0):
        self.n, self.m, self.C, self.H = int(n), int(m), int(Ccat), int(horizon)
        if KERNEL not in ("Matern", "RBF"):
            raise GpirtError(_lib.ERR_INVALID, f"KERNEL {KERNEL!r}: the reference only knows 'Matern' and 'RBF'")
        cfg = Config(n=self.n, m_total=self.m, horizon=self.H, C=self.C, constant_IRF=0, p_beta=3, theta_os=theta_os, theta_ls=theta_ls,
                     kernel=0 if KERNEL == "Matern" else 1, device=device, seed=seed, chain=chain, rank=0, world=1)
        self._L = _lib.lib()
        h = C.c_void_p()
        rc = self._L.gpirt_panel_create(C.byref(cfg), C.byref(h))
        self._h = h if h.value else None
        if rc != _lib.OK:
            raise GpirtError(rc, self._L.gpirt_last_error().decode())
        self.regime = "RDM" if (self.H > 1 and theta_ls <= 0.1) else ("GP" if self.H > 1 else "H1")
        self.S = []
        for k in range(self.H):
            sh = C.c_void_p()
            check(self._L.gpirt_panel_session(self._h, k, C.byref(sh)))
            self.S.append(Sampler._borrowed(sh, self.n, self.m, self.C))

    def close(self):
        if getattr(self, "_h", None):
            for s in self.S:
                s.close()
            self._L.gpirt_panel_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- cubes in, cubes out ----
    def set_data(self, y):
        y = np.asarray(y)
        if y.shape != (self.n, self.m, self.H):
            raise GpirtError(_lib.ERR_SHAPE, f"y has shape {y.shape}, expected {(self.n, self.m, self.H)}")
        for k, s in enumerate(self.S):
            s.set_data(y[:, :, k])

    def set_priors(self, beta_prior_sds=None, theta_prior_sds=None):
        for s in self.S:
            s.set_priors(beta_prior_sds, theta_prior_sds)

    def set_state(self, theta=None, beta=None, f=None, thresholds=None, fstar=None):
        pick = lambda a, k, nd: None if a is None else np.take(np.asarray(a, dtype=np.float64), k, axis=nd)
        for k, s in enumerate(self.S):
            s.set_state(theta=pick(theta, k, 1), beta=pick(beta, k, 2), f=pick(f, k, 2), thresholds=pick(thresholds, k, 2), fstar=pick(fstar, k, 2))

    def get_state(self, fstar=False):
        st = [s.get_state(fstar=fstar) for s in self.S]
        out = dict(theta=np.stack([x["theta"] for x in st], axis=1), beta=np.stack([x["beta"] for x in st], axis=2),
                   f=np.stack([x["f"] for x in st], axis=2), thresholds=np.stack([x["thresholds"] for x in st], axis=2))
        if fstar:
            out["fstar"] = np.stack([x["fstar"] for x in st], axis=2)
        return out

    def set_iter(self, it):
        for s in self.S:
            s.set_iter(it)

    def L_time(self):
        out = np.empty((self.H, self.H), order="F")
        check(self._L.gpirt_panel_get_L_time(self._h, _p(out)))
        return out

    # ---- blocks ----
    def draw_theta(self, z=None, u=None):
        """Philox streams when z and u are None. Parity mode: RDM takes lists of H arrays (z[h]: (1, n), u[h]: (ustride, n)); GP takes
        z (>= H, n) and u (ustride, n)."""
        if z is None and u is None:
            check(self._L.gpirt_panel_draw_theta(self._h, None))
            return
        if self.regime == "GP":
            s, keep = _streams(z, u)
            check(self._L.gpirt_panel_draw_theta(self._h, C.byref(s)))
        else:
            arr = (Streams * self.H)()
            keep = []
            for k in range(self.H):
                sk, kk = _streams(None if z is None else z[k], None if u is None else u[k])
                keep.append(kk)
                for fld, _ in Streams._fields_:
                    setattr(arr[k], fld, getattr(sk, fld))
            check(self._L.gpirt_panel_draw_theta(self._h, arr))
        self.sync()  # the staged host arrays must outlive the copies
        del keep

    def sweep(self, k=1):
        check(self._L.gpirt_panel_sweep(self._h, int(k)))

    def total_ll(self):
        tot = C.c_double()
        check(self._L.gpirt_panel_total_ll(self._h, C.byref(tot)))
        return tot.value

    def sync(self):
        check(self._L.gpirt_panel_sync(self._h))

    def diag_theta(self):
        """(evals, margin): RDM -> (H, n) arrays, one row per session; GP -> (n,) arrays of the joint step."""
        if self.regime == "GP":
            return self.S[0].diag("theta")[:2]
        d = [s.diag("theta") for s in self.S]
        return np.stack([x[0] for x in d]), np.stack([x[1] for x in d])
