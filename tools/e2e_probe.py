import os, sys, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import gpirt_b200
from bench import synth, N_RESP, M_PER_GPU, C_CAT
n,m,C=N_RESP,M_PER_GPU,C_CAT
P=synth(n,m,C,20261104+3)
y_pin=torch.empty((m,n),dtype=torch.float64,pin_memory=True); y_host=y_pin.numpy().T
y_host[...]=np.where(P["y"]==0,np.nan,P["y"].astype(np.float64))
for rep in range(6):
    for k in (20, 100):
        t0=time.perf_counter()
        res=gpirt_b200.gpirtMCMC(y_host,sample_iterations=k,burn_iterations=0,THIN=1,CHAIN=1,beta_prior_sds=P["beta_prior_sds"],theta_prior_sds=P["theta_prior_sds"],theta_init=P["theta0"],thresholds=P["thr0"],SEED=20261104,beta_init=P["beta0"])
        dt=time.perf_counter()-t0
        print(rep,k,f"{dt:.3f}s  {k/dt:.1f} sweeps/s", flush=True)
