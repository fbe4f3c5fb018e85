"""Bit-reproducibility stress of the Cholesky paths on small / ragged sizes: python tools/chol_determinism.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200 as g

rng = np.random.default_rng(1)
bad = 0
for n in (60, 100, 180, 333, 1001):
    x = -5.0 + np.round((np.clip(rng.standard_normal(n), -5, 5) + 5.0) / 0.01) * 0.01
    A = g.K(x, x, np.array([3.0, 3.0, 3.0])) + 1e-6 * np.eye(n)
    L0 = g.chol(A)
    diff = 0
    for rep in range(30):
        L = g.chol(A)
        if not np.array_equal(L, L0):
            diff += 1
    back = np.max(np.abs(L0 @ L0.T - A)) / np.max(np.abs(A))
    print(f"n={n}: {diff} of 30 repetitions differ from the first; backward {back:.2e}", flush=True)
    bad += diff
print("DETERMINISTIC" if bad == 0 else "NOT DETERMINISTIC")
