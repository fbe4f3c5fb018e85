"""Forward error of the f* draw on the reference's own datasets: CUDA (block-inverse solver, substitution solver) and the CPU oracle against
an 80-bit long-double evaluation of the same formulas (src/draw-fstar.cpp:26-55) from the same state and the same z.
   python tools/fstar_accuracy.py senate116|sdo          (GPU box; reads tests/golden/<name>.npz)
The error is reported norm-wise per item column, max |x - truth| / max(1, max |truth|), the measure the parity tests use."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gpirt_b200 as g  # noqa: E402
from oracle import orc  # noqa: E402
from tests.test_datasets import colrel, load, oracle_run  # noqa: E402

LD = np.longdouble


def chol_ld(A):
    A = A.astype(LD).copy()
    n = A.shape[0]
    for j in range(n):
        A[j, j] = np.sqrt(A[j, j] - A[j, :j] @ A[j, :j])
        A[j + 1:, j] = (A[j + 1:, j] - A[j + 1:, :j] @ A[j, :j]) / A[j, j]
    return np.tril(A)


def solve_lower_ld(L, B):
    X = B.astype(LD).copy()
    for i in range(L.shape[0]):
        X[i] = (X[i] - L[i, :i] @ X[:i]) / L[i, i]
    return X


def K_ld(x1, x2, s):
    x1, x2, s = x1.astype(LD)[:, None], x2.astype(LD)[None, :], s.astype(LD)
    return np.exp(LD(-0.5) * (x1 - x2) ** 2) + x1 * s[1] ** 2 * x2 + s[0] ** 2 + (x1 * s[2] * x2) ** 2


def fstar_truth(theta, f, beta, sds, z):
    grid = (LD(-5.0) + np.arange(1001) * 0.01).astype(np.float64)  # regspace(-5, 0.01, 5) as doubles
    n = theta.size
    L = chol_ld(K_ld(theta, theta, sds) + LD(1e-6) * np.eye(n, dtype=LD))
    tmp = solve_lower_ld(L, K_ld(theta, grid, sds))
    Kp = K_ld(grid, grid, sds) - tmp.T @ tmp + LD(1e-6) * np.eye(1001, dtype=LD)
    Lp = chol_ld(Kp)
    alpha = solve_lower_ld(L, f)  # L^-1 f ; mean = tmp^T alpha = k*^T L^-T L^-1 f
    X = np.stack([np.ones(1001), grid, grid * grid], axis=1).astype(LD)
    return (tmp.T @ alpha + X @ beta.astype(LD) + Lp @ z.astype(LD)).astype(np.float64)


def main(name):
    G = load(name)
    n, m, C = int(G["n"]), int(G["m"]), int(G["C"])
    S = None
    for it, rec, s, start in oracle_run(G):
        if S is None:
            S = g.Sampler(n, m, C)
            S.set_data(start["y"])
            S.set_priors(start["bps"], start["tps"])
            S.set_state(theta=G["theta_in"], beta=start["beta0"], f=start["f0"], thresholds=start["thr0"], fstar=start["fstar0"])
        S.draw_f(z=rec["f"][0], u=rec["f"][1])
        st = S.get_state()
        kk = min(m, 24)
        truth = fstar_truth(st["theta"], st["f"][:, :kk], st["beta"][:, :kk], start["bps"][:, 0], rec["fstar"][0][:, :kk])
        out = {}
        for solver in (1, 0):
            S.set_solver(solver)
            S.draw_fstar(z=rec["fstar"][0])
            out[solver] = S.get_state(fstar=True)["fstar"][:, :kk]
        S.set_solver(1)
        S.draw_fstar(z=rec["fstar"][0])
        # the oracle's f* of this sweep was drawn from ITS f (differs from the CUDA f by ~1e-8), so it is compared on its own inputs
        print(f"{name} sweep {it}: |f*| max {np.abs(truth).max():.1f}  cuda block-inverse {colrel(out[1], truth):.2e}  cuda substitution {colrel(out[0], truth):.2e}"
              f"  block-inverse vs substitution {colrel(out[1], out[0]):.2e}  oracle vs cuda {colrel(s['fstar'][:, :kk], out[1]):.2e}")
        S.draw_theta(z=rec["theta"][0], u=rec["theta"][1])
        S.gather_f()
        S.draw_beta(z=rec["beta"][0], u=rec["beta"][1])
        S.update_cholesky_cache()
        S.draw_threshold(z=rec["thr"][0], u=rec["thr"][1])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "senate116")
