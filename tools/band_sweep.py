"""C3 sweeps timed with CUDA events (wall of 30 sweeps after 5) plus the gemm_LZ / gemm_FtW group times of a profiled pass, for the GEMM tile order
selected by the environment (GPIRT_GEMM_BAND_MB: 0 = one band, the row-slow order; default 24)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
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
best = 1e9
for rep in range(3):
    t0 = time.perf_counter()
    S.sweep(30)
    S.sync()
    best = min(best, (time.perf_counter() - t0) / 30 * 1e3)
S.profile(True)
S.sweep(1)
S.profile_reset()
S.sweep(4)
sec = S.profile_read()
S.profile(False)
print(f"band_mb={os.environ.get('GPIRT_GEMM_BAND_MB', 'default')}: {best:.3f} ms/sweep ({1e3 / best:.1f}/s)  " +
      " ".join(f"{k}={sec[k][0] / 4:.4f}" for k in ("ess_f", "ess_beta", "ess_thr", "ess_theta", "gemm_LZ", "gemm_FtW", "gemm_ZLpost", "gemm_tmp", "gemm_W", "gemm_Kpost", "trtri", "chol_K")))
