"""gpirtMCMC() end to end at C3 for k = 1, 10, 20, 40 kept sweeps: wall clock = a + b k. a is what a call costs besides its sweeps (handle, fp64 y upload
from pinned memory, initial draws, read-back, Python marshalling); the C call alone is timed next to the whole Python call."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200  # noqa: E402
from gpirt_b200 import _lib  # noqa: E402
from bench import C_CAT, M_PER_GPU, N_RESP, host_responses, synth  # noqa: E402

n, m, C = N_RESP, M_PER_GPU, C_CAT
P = synth(n, m, C, 20261104 + 3)
y_host = host_responses(torch, P["y"], pinned=True)
L = _lib.lib()
real = L.gpirt_mcmc
inner = []


class Timed:
    def __call__(self, *a):
        t0 = time.perf_counter()
        rc = real(*a)
        inner.append(time.perf_counter() - t0)
        return rc


L.gpirt_mcmc = Timed()
rows = []
for k in (1, 10, 20, 40, 20):
    tot, inn = [], []
    for _ in range(5):
        inner.clear()
        t0 = time.perf_counter()
        gpirt_b200.gpirtMCMC(y_host, sample_iterations=k, burn_iterations=0, THIN=1, CHAIN=1, beta_prior_sds=P["beta_prior_sds"], theta_prior_sds=P["theta_prior_sds"],
                             theta_init=P["theta0"], thresholds=P["thr0"], SEED=20261104, beta_init=P["beta0"])
        tot.append(time.perf_counter() - t0)
        inn.append(inner[-1])
    rows.append((k, np.median(tot) * 1e3, np.median(inn) * 1e3))
    print(f"k = {k:3d}: whole call {rows[-1][1]:8.2f} ms, C call {rows[-1][2]:8.2f} ms, Python around it {rows[-1][1] - rows[-1][2]:6.2f} ms", flush=True)
ks = np.array([r[0] for r in rows[:4]], dtype=float)
for col, name in ((1, "whole call"), (2, "C call")):
    b, a = np.polyfit(ks, np.array([r[col] for r in rows[:4]]), 1)
    print(f"{name}: {a:.2f} ms + {b:.3f} ms per sweep")
