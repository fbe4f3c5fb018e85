"""Two item shards of a C3-sized problem on ONE GPU (LocalShards): the command wrapped by ncu for the item-sharded theta kernels."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import C_CAT, M_PER_GPU, N_RESP, synth  # noqa: E402
from gpirt_b200.distributed import LocalShards  # noqa: E402

w, k = int(sys.argv[1]), int(sys.argv[2])
n, m, C = N_RESP, 2 * M_PER_GPU, C_CAT
P = synth(n, m, C, 20261104 + 3)
S = LocalShards(n, m, C, devices=[0, 0], seed=20261104)
S.set_data(P["y"])
S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
S.init_draw(draw_beta=False, draw_f=True)
S.sweep(w)
S.sync()
S.sweep(k)
S.sync()
ev = S.diag("theta")[0]
print("ll", S.total_ll(), "theta rounds in the last sweep", S.theta_rounds, "mean theta evaluations", ev.mean())
