"""constant_IRF == 1: one item response function shared by all sessions (src/draw-f.cpp:84-139, src/draw-fstar.cpp:58-125,
src/draw_threshold.cpp:5-96,181-205, initialisation src/gpirtMCMC.cpp:164-200), composed from the handles the library already has:

  P   gpirt_panel            the H sessions: joint theta step, f <- f* gather, draw_beta, total log-likelihood — these take no constant_IRF
                             argument in the reference and run unchanged
  T   gpirt_sampler, n*H     the STACKED sampler: row h*n + i is respondent i in session h. draw_f = one ESS per item over the stacked column
                             with nu = chol(K(theta_stacked) + 1e-6 I) z and the rows' own means (gpirt_set_mean_rows: mu_h = X_h beta_h differs
                             by session); draw_threshold = one ESS per item over all sessions' cells, session 0's thresholds
  I   gpirt_sampler, 100     the inducing-point sampler: src/draw-fstar.cpp:84-117 is the ordinary f* draw (:26-55) with theta -> the 100
                             inducing points linspace(min theta, max theta) and f -> the stacked f interpolated onto them (gpirt_interp1)

Every likelihood / ESS / linear-algebra step runs in the CUDA kernels; this module only moves state between the handles (host copies — the
mode is functional, not tuned) and forms mu_h = X_h beta_h and the linspace. The same composition as the test oracle's ConstantPanel, which is pinned to the
reference's own constant_IRF == 1 runs. The C entry point gpirt_mcmc still refuses constant_IRF == 1; gpirt_b200.gpirtMCMC routes here.
"""
import numpy as np

from . import _lib
from ._lib import NGRID, GpirtError
from .api import interp1
from .panel import Panel
from .sampler import Sampler

N_INDUCED = 100  # src/draw-fstar.cpp:73


class ConstantPanel:
    def __init__(self, n, m, Ccat, horizon, *, theta_os=1.0, theta_ls=10.0, KERNEL="Matern", seed=0, chain=0, device=0):
        self.n, self.m, self.C, self.H, self.device = int(n), int(m), int(Ccat), int(horizon), device
        self.P = Panel(n, m, Ccat, horizon, theta_os=theta_os, theta_ls=theta_ls, KERNEL=KERNEL, seed=seed, chain=chain, device=device)
        self.S = self.P.S
        self.T = Sampler(self.n * self.H, m, Ccat, seed=seed, chain=chain, device=device)
        self.I = Sampler(N_INDUCED, m, Ccat, seed=seed, chain=chain, device=device)
        self.mu = None  # n x m x H: what src/gpirtMCMC.cpp keeps in `mu`
        self.iter = 0

    def close(self):
        for h in (self.T, self.I, self.P):
            h.close()

    # ---- data / state ----
    def set_data(self, y):
        y = np.asarray(y, dtype=np.float64)
        self.P.set_data(y)
        self.T.set_data(np.concatenate([y[:, :, h] for h in range(self.H)], axis=0))  # row h*n + i, src/draw-f.cpp:94-103
        self.I.set_data(np.zeros((N_INDUCED, self.m), dtype=np.int8))  # the inducing sampler only draws f*: no responses

    def set_priors(self, beta_prior_sds, theta_prior_sds=None):
        tps = np.zeros((2, self.n)) if theta_prior_sds is None else np.asarray(theta_prior_sds, dtype=np.float64)
        self.P.set_priors(beta_prior_sds, tps)
        self.T.set_priors(beta_prior_sds, np.tile(tps, (1, self.H)))
        self.I.set_priors(beta_prior_sds, np.zeros((2, N_INDUCED)))

    @staticmethod
    def _X_beta(theta, beta):  # mu = X beta, X = [1, theta, theta^2], accumulated left to right (src/gpirtMCMC.cpp:308)
        t = np.asarray(theta)[:, None]
        return (beta[0][None, :] + t * beta[1][None, :]) + (t * t) * beta[2][None, :]

    def init_constant(self, theta, beta0, f0, thresholds):
        """The state the constant branch of the initialisation leaves (src/gpirtMCMC.cpp:164-200): beta for session 0 only (the other slices are
        zero), the same f0 columns in every session, every session's mu = X_h beta_0."""
        theta = np.clip(np.asarray(theta, dtype=np.float64), -5.0, 5.0)
        thr = np.asarray(thresholds, dtype=np.float64)
        for h, S in enumerate(self.S):
            S.set_state(theta=theta[:, h], beta=beta0 if h == 0 else np.zeros_like(beta0), f=f0, thresholds=thr[:, :, h])
        self.mu = np.stack([self._X_beta(theta[:, h], np.asarray(beta0)) for h in range(self.H)], axis=2)

    def set_iter(self, it):
        self.iter = it
        for h in (self.P, self.T, self.I):
            h.set_iter(it)

    def get_state(self, fstar=True):
        return self.P.get_state(fstar=fstar)

    def _stacked(self):
        st = [S.get_state() for S in self.S]
        theta = np.concatenate([s["theta"] for s in st])
        f = np.concatenate([s["f"] for s in st], axis=0)
        mu = np.concatenate([self.mu[:, :, h] for h in range(self.H)], axis=0)
        return st, theta, f, mu

    # ---- blocks ----
    def draw_f(self, z=None, u=None):
        st, theta, f, mu = self._stacked()
        self.T.set_state(theta=theta, beta=st[0]["beta"], f=f, thresholds=st[0]["thresholds"])
        self.T.set_mean_rows(mu)
        self.T.draw_f(z=z, u=u)
        fn = self.T.get_state()["f"]
        for h, S in enumerate(self.S):
            S.set_state(f=fn[h * self.n:(h + 1) * self.n])

    def draw_fstar(self, z=None):
        st, theta, f, _ = self._stacked()
        lo, hi = float(theta.min()), float(theta.max())
        tind = lo + np.arange(N_INDUCED, dtype=np.float64) * ((hi - lo) / (N_INDUCED - 1))  # arma::linspace: start + i*delta, last = end
        tind[-1] = hi
        find = interp1(theta, f, tind, device=self.device)  # src/draw-fstar.cpp:77-82
        self.I.set_state(theta=tind, beta=st[0]["beta"], f=find)
        self.I.draw_fstar(z=z)
        fs = self.I.get_state(fstar=True)["fstar"]
        for S in self.S:  # one f* for all sessions (:121-124)
            S.set_state(fstar=fs)

    def draw_theta(self, z=None, u=None):
        self.P.draw_theta(z=z, u=u)

    def gather_f(self):
        for S in self.S:
            S.gather_f()

    def draw_beta(self, z=None, u=None):
        for h, S in enumerate(self.S):
            S.draw_beta(z=None if z is None else z[h], u=None if u is None else u[h])
        st = [S.get_state() for S in self.S]
        self.mu = np.stack([self._X_beta(s["theta"], s["beta"]) for s in st], axis=2)  # src/gpirtMCMC.cpp:307-311

    def draw_threshold(self, z=None, u=None):
        st, theta, f, mu = self._stacked()
        self.T.set_state(theta=theta, f=f, thresholds=st[0]["thresholds"])
        self.T.set_mean_rows(mu)
        self.T.draw_threshold(z=z, u=u)
        thr = self.T.get_state()["thresholds"]
        for S in self.S:  # src/draw_threshold.cpp:199-203
            S.set_state(thresholds=thr)

    def total_ll(self):
        acc = 0.0  # one running sum in session order, src/gpirtMCMC.cpp:324-331 (not the builtin sum(): Python >= 3.12 compensates float sums,
        for S in self.S:  # which is not the reference's left-to-right accumulation and differs from it in the last bit now and then)
            acc += S.ll_bar()[1]
        return acc

    def sweep(self):
        """One sweep (src/gpirtMCMC.cpp:267-331), Philox streams."""
        self.draw_f()
        self.draw_fstar()
        self.draw_theta()
        self.gather_f()
        self.draw_beta()
        for S in self.S:
            S.update_cholesky_cache()
        self.draw_threshold()
        self.set_iter(self.iter + 1)
