"""Runs the C3 chain several times (overlap on/off, screening on/off) and reports whether the states agree bit for bit."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200  # noqa: E402
from bench import C_CAT, M_PER_GPU, N_RESP, synth  # noqa: E402

sweeps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n, m, C = N_RESP, M_PER_GPU, C_CAT
P = synth(n, m, C, 20261104 + 3)


def run(overlap, screen, solver=1):
    S = gpirt_b200.Sampler(n, m, C, seed=20261104)
    S.set_overlap(overlap); S.set_screen(screen); S.set_solver(solver)
    S.set_data(P["y"]); S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True)
    out = []
    for k in range(sweeps):
        S.sweep(1)
        out.append((S.total_ll(), S.get_state(f=False, thresholds=False, beta=False)["theta"].copy()))
    S.close()
    return out


ref = run(0, 0)
for name, cfg in (("serial again", (0, 0)), ("overlap", (1, 0)), ("overlap again", (1, 0)), ("screen", (0, 1)), ("overlap+screen", (1, 1)), ("overlap+screen again", (1, 1))):
    got = run(*cfg)
    first = next((k for k in range(sweeps) if got[k][0] != ref[k][0] or not np.array_equal(got[k][1], ref[k][1])), None)
    print(f"{name:22s} first differing sweep: {first}   ll[-1] {got[-1][0]!r} vs {ref[-1][0]!r}")
# same, but without a host sync between sweeps (the bench's pattern)
def run_nosync(overlap, screen):
    S = gpirt_b200.Sampler(n, m, C, seed=20261104)
    S.set_overlap(overlap); S.set_screen(screen)
    S.set_data(P["y"]); S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True)
    S.sweep(sweeps); S.sync()
    ll = S.total_ll(); S.close()
    return ll
for cfg in ((0, 0), (1, 0), (1, 0), (1, 1), (1, 1)):
    print("no-sync", cfg, repr(run_nosync(*cfg)), "ref", repr(ref[-1][0]))
