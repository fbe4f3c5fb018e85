"""One GPU's share of BASELINE config C4 (n = 10 000 respondents, 50 000 items over 8 GPUs -> 6250 items here, C = 7): sweeps run, stay finite, and
two identically seeded chains agree bit for bit. Prints ms per sweep and the per-group device times."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gpirt_b200
from tests.problems import make_problem

n, m, C = 10000, int(sys.argv[1]) if len(sys.argv) > 1 else 6250, 7
P = make_problem(n, m, C, seed=20261104 + 4, missing=0.05, dtype="i8")
def chain(k):
    S = gpirt_b200.Sampler(n, m, C, seed=1234)
    S.set_data(P["y"]); S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True)
    S.sweep(2); S.sync()
    t0 = time.perf_counter(); S.sweep(k); S.sync(); dt = (time.perf_counter() - t0) / k
    return S, dt
S, dt = chain(5)
ll = S.total_ll(); th = S.get_state(f=False)["theta"]
print(f"C4 shard n={n} m={m} C={C}: {dt*1e3:.2f} ms/sweep, ll {ll:.6f}, mean evals", {b: float(S.diag(b)[0].mean()) for b in ("f", "theta", "beta", "thr")})
S.profile(1); S.profile_reset(); S.sweep(2); sec = S.profile_read(); S.profile(0)
print("  " + " ".join(f"{k}={v[0]/2:.2f}" for k, v in sorted(sec.items(), key=lambda kv: -kv[1][0])[:14]))
S.close()
S2, _ = chain(5)
assert np.isfinite(ll) and S2.total_ll() == ll and np.array_equal(S2.get_state(f=False)["theta"], th), "chains differ"
print("deterministic: ok")
