"""Sets the C3 workload up and runs W + K plain sweeps — the command wrapped by ncu (launch list / --set full)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200  # noqa: E402
from bench import C_CAT, M_PER_GPU, N_RESP, synth  # noqa: E402

w, k = int(sys.argv[1]), int(sys.argv[2])
n, m, C = N_RESP, (int(sys.argv[3]) if len(sys.argv) > 3 else M_PER_GPU), C_CAT
P = synth(n, m, C, 20261104 + 3)
S = gpirt_b200.Sampler(n, m, C, seed=20261104)
S.set_data(P["y"])
S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
S.init_draw(draw_beta=False, draw_f=True)
S.sweep(w)
S.sync()
S.sweep(k)
S.sync()
print("ll", S.total_ll(), "mean evaluations per unit in the last sweep:", {b: round(float(S.diag(b)[0].mean()), 2) for b in ("f", "theta", "beta", "thr")})
