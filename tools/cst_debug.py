"""Which draws differ between gpirtMCMC(constant_IRF=1) through the C entry point and the stepwise ConstantPanel loop."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpirt_b200 as g
from gpirt_b200.api import _as_cube, _gpirtMCMC_constant_irf
from tests.problems import make_problem

n, m, C, H = 60, 9, 4, 3
ls = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
Ps = [make_problem(n, m, C, seed=300 + h, missing=0.1) for h in range(H)]
y = np.asfortranarray(np.stack([P["y"] for P in Ps], axis=2))
th0 = np.asfortranarray(np.stack([Ps[0]["theta0"]] * H, axis=1))
kw = dict(theta_ls=ls, theta_os=1.2, theta_init=th0, thresholds=Ps[0]["thr0"], SEED=9, constant_IRF=1, store_f=True, store_fstar=True)
for rep in range(2):
    a = g.gpirtMCMC(y, 4, 2, THIN=2, **kw)[0]
    b = _gpirtMCMC_constant_irf(_as_cube(y), 4, 2, 2, 1, None, None, None, None, 1.2, ls, "Matern", th0, Ps[0]["thr0"], 9, True, True, 0, None, 0)[0]
    print("rep", rep, "theta", np.max(np.abs(a["theta"] - b["theta"])), "ll", a["ll"] - b["ll"])
    for k in range(2):
        for key in ("beta", "threshold", "f", "fstar"):
            x, z = a[key][k], b[key][k]
            fin = np.isfinite(x) & np.isfinite(z)
            print("  draw", k, key, "max abs diff", np.max(np.abs(x[fin] - z[fin])), "of scale", np.max(np.abs(z[fin])))
