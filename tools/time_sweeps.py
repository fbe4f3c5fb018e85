"""ms per C3 sweep (wall clock around K sweeps bracketed by device syncs, best of R repeats) — for quick A/B runs of env knobs."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200  # noqa: E402
from bench import C_CAT, M_PER_GPU, N_RESP, synth  # noqa: E402

k = int(sys.argv[1]) if len(sys.argv) > 1 else 40
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
n, m, C = N_RESP, M_PER_GPU, C_CAT
P = synth(n, m, C, 20261104 + 3)
S = gpirt_b200.Sampler(n, m, C, seed=20261104)
S.set_data(P["y"])
S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
S.init_draw(draw_beta=False, draw_f=True)
S.sweep(5)
S.sync()
ts = []
for _ in range(reps):
    t0 = time.perf_counter()
    S.sweep(k)
    S.sync()
    ts.append((time.perf_counter() - t0) / k * 1e3)
print("ms/sweep best %.3f median %.3f  ll %.6f" % (min(ts), sorted(ts)[len(ts) // 2], S.total_ll()), {e: os.environ[e] for e in os.environ if e.startswith("GPIRT_")})
