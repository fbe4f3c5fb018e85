"""Where the end-to-end gpirtMCMC() call spends its wall clock at C3: the same steps through the Sampler API, timed one by one."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import gpirt_b200
from bench import synth, N_RESP, M_PER_GPU, C_CAT
n, m, C = N_RESP, M_PER_GPU, C_CAT
P = synth(n, m, C, 20261104 + 3)
y_pin = torch.empty((m, n), dtype=torch.float64, pin_memory=True); y_host = y_pin.numpy().T
y_host[...] = np.where(P["y"] == 0, np.nan, P["y"].astype(np.float64))
K = 20
for rep in range(4):
    t = [time.perf_counter()]
    def lap(): t.append(time.perf_counter())
    S = gpirt_b200.Sampler(n, m, C, seed=20261104); lap()
    S.set_data(y_host); S.sync(); lap()
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"]); S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"]); S.sync(); lap()
    S.init_draw(draw_beta=False, draw_f=True); S.sync(); lap()
    for k in range(K):
        S.sweep(1); st = S.get_state(); ll = S.total_ll()
    lap()
    S.sweep(K); S.sync(); lap()
    del S; lap()
    names = ["create", "set_data(f64 y)", "priors+state", "init_draw", f"{K} x (sweep+get_state+ll)", f"sweep({K}) no readback", "destroy"]
    print(rep, "  ".join(f"{nm} {1e3 * (b - a):.2f} ms" for nm, a, b in zip(names, t[:-1], t[1:])), flush=True)
    t0 = time.perf_counter()
    res = gpirt_b200.gpirtMCMC(y_host, sample_iterations=K, burn_iterations=0, THIN=1, CHAIN=1, beta_prior_sds=P["beta_prior_sds"], theta_prior_sds=P["theta_prior_sds"],
                               theta_init=P["theta0"], thresholds=P["thr0"], SEED=20261104, beta_init=P["beta0"])
    print(rep, f"gpirtMCMC({K}) {1e3 * (time.perf_counter() - t0):.2f} ms", flush=True)
