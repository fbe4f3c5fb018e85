"""Per-section elapsed times of the OVERLAPPED schedule at C3 (profiling level 2: waits and SM sharing included)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200
from bench import C_CAT, M_PER_GPU, N_RESP, synth
n, m, C = N_RESP, M_PER_GPU, C_CAT
P = synth(n, m, C, 20261104 + 3)
S = gpirt_b200.Sampler(n, m, C, seed=20261104)
S.set_data(P["y"]); S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
S.init_draw(draw_beta=False, draw_f=True)
S.sweep(5); S.sync()
for level in (1, 2):
    S.profile(level); S.profile_reset(); S.sweep(5)
    sec = S.profile_read(); S.profile(0)
    tot = sum(v[0] for v in sec.values()) / 5
    print(f"level {level}: sum {tot:.3f} ms/sweep")
    print("  " + " ".join(f"{k}={v[0]/5:.3f}" for k, v in sorted(sec.items(), key=lambda kv: -kv[1][0])))
