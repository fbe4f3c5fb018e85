"""Seeded synthetic problems shared by the tests, smoke() and bench.py.

theta_true ~ N(0,1) clipped to [-5,5]; g_ij = a_j + b_j theta_i + 0.5 sin(c_j theta_i); thresholds qnorm(k/C);
y_ij = 1 + #{k: g_ij + eps_ij > tau_k}; a fraction `missing` of cells set to NaN; reference default priors.
"""
import numpy as np

from gpirt_b200.sampler import default_thresholds


def make_problem(n, m, C, seed=20261104, missing=0.05, on_grid=False, dtype="f64"):
    rng = np.random.default_rng(seed)
    theta_true = np.clip(rng.standard_normal(n), -5, 5)
    a = rng.standard_normal(m)
    b = rng.uniform(0.5, 2.0, m) * rng.choice([-1.0, 1.0], m)
    c = rng.uniform(0.5, 2.0, m)
    tau = default_thresholds(1, C)[0]
    y = np.empty((n, m), order="F", dtype=np.float64 if dtype == "f64" else np.int8)
    blk = max(1, (1 << 22) // max(n, 1))
    for j0 in range(0, m, blk):
        j1 = min(m, j0 + blk)
        g = a[None, j0:j1] + theta_true[:, None] * b[None, j0:j1] + 0.5 * np.sin(theta_true[:, None] * c[None, j0:j1])
        g = g + rng.standard_normal(g.shape)
        yy = 1 + (g[:, :, None] > tau[None, None, 1:C]).sum(axis=2)
        miss = rng.random(g.shape) < missing
        if dtype == "f64":
            yy = yy.astype(np.float64)
            yy[miss] = np.nan
        else:
            yy = yy.astype(np.int8)
            yy[miss] = 0
        y[:, j0:j1] = yy
    theta0 = np.clip(rng.standard_normal(n), -5, 5)
    if on_grid:  # the state after any sweep: theta lives on the 0.01 grid (src/draw-theta.cpp:165-169)
        theta0 = -5.0 + np.round((theta0 + 5.0) / 0.01) * 0.01
    beta0 = np.asfortranarray(rng.standard_normal((3, m)) * 3.0)
    thr0 = default_thresholds(m, C)
    return dict(n=n, m=m, C=C, y=y, theta0=theta0, beta0=beta0, thr0=thr0,
                beta_prior_sds=np.full((3, m), 3.0, order="F"), theta_prior_sds=np.zeros((2, n), order="F"), rng=rng)


def y_to_i8(y):
    out = np.where(np.isfinite(y), y, 0).astype(np.int8)
    return np.asfortranarray(out)


def uniform_streams(rng, units, stride=64):
    return np.asfortranarray(rng.random((stride, units)))
