"""C3 on one GPU: the overlapped schedule's own section timeline (profiling level 2: section times include waiting and sharing the SMs), next to the
serial order's (level 1). The wait_* sections are the main stream's stalls on the theta-only chain."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200  # noqa: E402
from bench import C_CAT, M_PER_GPU, N_RESP, synth  # noqa: E402

n, m, C = N_RESP, M_PER_GPU, C_CAT
P = synth(n, m, C, 20261104 + 3)
S = gpirt_b200.Sampler(n, m, C, seed=20261104)
S.set_data(P["y"])
S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
S.init_draw(draw_beta=False, draw_f=True)
S.sweep(5)
S.sync()
for level in (2, 1):
    S.profile(level)
    S.sweep(1)
    S.profile_reset()
    S.sweep(6)
    sec = S.profile_read()
    S.profile(0)
    main = ("gen_Z", "wait_L", "gemm_LZ", "ess_f", "wait_W", "gemm_FtW", "wait_shared", "gemm_ZLpost", "add_mustar", "theta_table", "ess_theta", "ess_beta",
            "ess_thr", "item_ll", "gather_f")
    side = ("K_build", "chol_K", "trtri", "gemm_tmp", "gemm_W", "gemm_Kpost", "chol_Kpost")
    print(f"level {level}: main " + " ".join(f"{k}={sec[k][0] / 6:.3f}" for k in main if k in sec) + f" | sum {sum(sec[k][0] for k in main if k in sec) / 6:.3f}")
    print(f"         side " + " ".join(f"{k}={sec[k][0] / 6:.3f}" for k in side if k in sec) + f" | sum {sum(sec[k][0] for k in side if k in sec) / 6:.3f}")
    print("         other " + " ".join(f"{k}={v[0] / 6:.3f}" for k, v in sec.items() if k not in main and k not in side))

# the schedule as it ran: two consecutive sweeps, every section with its begin / end on the device clock
S.profile(2)
S.sweep(1)
S.profile_reset()
S.sweep(2)
tl = S.profile_timeline()
S.profile(0)
side_names = set(side)
print("timeline of two sweeps (ms after the first section's begin; M = main stream, S = side streams):")
for name, b, e in sorted(tl, key=lambda x: x[1]):
    print(f"  {'S' if name in side_names else 'M'} {name:12s} {b:7.3f} -> {e:7.3f}  ({e - b:.3f})")
