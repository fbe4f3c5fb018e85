"""Aggregate throughput of k chains batched on one GPU (gpirt_mcmc_chains' schedule: round-robin sweeps, per-chain streams) against one chain.
   python tools/chains_bench.py n m C chains sweeps          prints chain-sweeps/s for 1 and for `chains` chains (device time, CUDA events
on a separate timing stream are not needed: the wall clock brackets a full device synchronisation on both sides, after warm-up)."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import gpirt_b200 as g  # noqa: E402
from tests.problems import make_problem  # noqa: E402


def run(P, k, sweeps, warm=3):
    n, m, C = P["n"], P["m"], P["C"]
    S = []
    for c in range(k):
        s = g.Sampler(n, m, C, seed=11, chain=c)
        s.set_data(P["y"])
        s.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        s.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
        s.set_iter(0xFFFFFFFF)
        s.init_draw(draw_beta=False, draw_f=True)
        s.set_iter(0)
        S.append(s)
    for _ in range(warm):
        for s in S:
            s.sweep(1)
    for s in S:
        s.sync()
    t0 = time.perf_counter()
    for _ in range(sweeps):
        for s in S:
            s.sweep(1)
    for s in S:
        s.sync()
    dt = time.perf_counter() - t0
    ll = [s.total_ll() for s in S]
    for s in S:
        s.close()
    return k * sweeps / dt, ll


if __name__ == "__main__":
    n, m, C, k, sweeps = (int(v) for v in sys.argv[1:6])
    P = make_problem(n, m, C, dtype="i8" if n * m > 5e7 else "f64")
    one, _ = run(P, 1, sweeps)
    many, ll = run(P, k, sweeps)
    print(f"n={n} m={m} C={C}: 1 chain {one:.1f} sweeps/s; {k} chains batched {many:.1f} chain-sweeps/s ({many / one:.2f}x); ll {np.round(ll[:3], 1)}")
