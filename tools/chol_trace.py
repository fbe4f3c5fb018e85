"""Per-phase clock64 trace of the persistent Cholesky: python tools/chol_trace.py [n] [ctas]"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200 as g
from gpirt_b200 import _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
ctas = int(sys.argv[2]) if len(sys.argv) > 2 else 128
rng = np.random.default_rng(0)
x = np.sort(rng.uniform(-3, 3, n))
A = np.asfortranarray(g.K(x, x, np.array([0.5, 0.3, 0.0])) + np.eye(n) * 0.5)
T = (n + 63) // 64
tr = np.zeros(16 * T + 2048, dtype=np.int64)
for rep in range(2):
    _lib.check(_lib.lib().gpirt_debug_chol_trace(0, A.ctypes.data_as(C.c_void_p), C.c_int64(n), ctas, tr.ctypes.data_as(C.c_void_p)))
crit = tr[:16 * T].reshape(T, 16)
names = ["top->diag ready", "potf2", "store+publish diag", "wait/issue sub", "sub tile landed", "solve", "store+publish sub", "wait/issue next diag", "next diag landed", "update"]
d = np.diff(crit[:-1, :11], axis=1)
print(f"n={n} T={T} ctas={ctas}: critical CTA, cycles per phase (median over steps; first / last step)")
for i, nm in enumerate(names):
    print(f"  {nm:24s} {np.median(d[:, i]):9.0f}   {d[0, i]:9.0f} {d[-1, i]:9.0f}")
step = np.diff(crit[:, 0])
print(f"  step total: median {np.median(step):.0f} cycles, sum {crit[-1, 3] - crit[0, 0]:.0f} cycles = {(crit[-1, 3] - crit[0, 0]) / 1.965e3:.1f} us at 1965 MHz")
w = tr[16 * T:].reshape(256, 8)
w = w[w[:, 0] > 0]
kinds = (w[:, 7] >> 30) & 3
for kind, nm in ((1, "S"), (2, "U")):
    q = w[kinds == kind]
    if len(q) == 0:
        continue
    dd = np.diff(q[:, :6], axis=1)
    print(f"worker 1, {len(q)} {nm} tasks, median cycles: waits+issue {np.median(dd[:, 0]):.0f}, tiles landed {np.median(dd[:, 1]):.0f}, compute {np.median(dd[:, 2]):.0f}, "
          f"store {np.median(dd[:, 3]):.0f}, publish {np.median(dd[:, 4]):.0f}; task total {np.median(q[:, 5] - q[:, 0]):.0f}")
