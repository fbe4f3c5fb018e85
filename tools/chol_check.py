"""GPU check + timing of the Cholesky paths (persistent single kernel vs per-panel launches): backward error, agreement with LAPACK,
not-positive-definite reporting, and the device time of chol(K) at the C3 and K_post sizes through a sampler handle's profiling sections.
  GPIRT_CHOL_PERSISTENT=0|1  GPIRT_CHOL_CTAS=n  python tools/chol_check.py [quick]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200 as g
from tests.problems import make_problem

quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
rng = np.random.default_rng(0)
worst = 0.0
for n in ([5, 64, 65, 130, 333] if quick else [5, 64, 65, 100, 128, 130, 333, 700, 1001, 1500, 2000, 3000]):
    x = np.sort(rng.uniform(-3, 3, n))
    A = g.K(x, x, np.array([0.5, 0.3, 0.0])) + np.eye(n) * 0.5
    L = g.chol(A)
    back = np.max(np.abs(L @ L.T - A)) / np.max(np.abs(A))
    fwd = np.max(np.abs(L - np.linalg.cholesky(A))) / np.max(np.abs(L))
    up = np.max(np.abs(np.triu(L, 1)))
    worst = max(worst, back)
    print(f"n={n:5d} backward {back:.2e} vs LAPACK {fwd:.2e} upper {up:.1e}", flush=True)
    assert back < 1e-14 and fwd < 1e-10 and up == 0.0
A = np.eye(70)
A[65, 65] = -1.0
try:
    g.chol(A)
    raise SystemExit("not-PD matrix was accepted")
except g.GpirtError as e:
    assert e.code == 4 and "66" in str(e), str(e)
    print("not-PD reported:", e)
if quick:
    raise SystemExit(0)
for n in (2000, 1001):
    P = make_problem(n, 8, 5, seed=3, on_grid=True)
    S = g.Sampler(n, 8, 5)
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.update_cholesky_cache()
    L0 = S.get_L()
    Kmat = g.K(P["theta0"], P["theta0"], np.array([3.0, 3.0, 3.0])) + 1e-6 * np.eye(n)
    print(f"default priors n={n}: backward {np.max(np.abs(L0 @ L0.T - Kmat)) / np.max(np.abs(Kmat)):.2e}")
    S.profile(True)
    S.profile_reset()
    reps = 20
    for _ in range(reps):
        S.set_state(theta=P["theta0"])
        S.update_cholesky_cache()
    S.sync()
    sec = S.profile_read()
    S.profile(False)
    assert np.array_equal(S.get_L(), L0), "factor changed between runs"
    print(f"n={n}: chol_K {sec['chol_K'][0] / reps:.4f} ms per factorisation ({int(sec['chol_K'][1] / reps)} timed section(s) each), "
          f"persistent={os.environ.get('GPIRT_CHOL_PERSISTENT', '1')} ctas={os.environ.get('GPIRT_CHOL_CTAS', 'default')}")
    S.close()
