"""Fits the fixed-coefficient approximations used by csrc/probit.cuh (run once; coefficients are pasted into the header).
Q(a) = Phi(-a), a >= 0:  Q(a) = exp(-a^2/2) * P(s),  s = (a - c)/(a + c)."""
import mpmath as mp, numpy as np, sys
mp.mp.dps = 50
A = 8.6
def R(a):
    a = mp.mpf(a)
    return mp.ncdf(-a) * mp.exp(a * a / 2)
def fit(c, deg, npts=400):
    smin, smax = -1.0, (A - c) / (A + c)
    # Chebyshev nodes on [smin, smax]
    k = np.arange(npts)
    x = np.cos(np.pi * (k + 0.5) / npts)
    s = 0.5 * (smax - smin) * x + 0.5 * (smax + smin)
    a = c * (1 + s) / (1 - s)
    y = [R(ai) for ai in a]
    # chebyshev coefficients via discrete orthogonality in mp
    coefs = []
    for j in range(deg + 1):
        acc = mp.mpf(0)
        for kk in range(npts):
            acc += y[kk] * mp.cos(mp.pi * j * (kk + 0.5) / npts)
        coefs.append(acc * 2 / npts)
    coefs[0] /= 2
    return coefs, (smin, smax)
def cheb_eval(coefs, x):
    b1 = b2 = mp.mpf(0)
    for cj in reversed(coefs[1:]):
        b1, b2 = 2 * x * b1 - b2 + cj, b1
    return x * b1 - b2 + coefs[0]
for c in (2.0, 3.0, 4.0):
    for deg in (16, 18, 20, 22, 24):
        coefs, (smin, smax) = fit(c, deg)
        worst = 0
        for a in np.linspace(0, A, 1500):
            s = (a - c) / (a + c)
            x = (2 * s - (smax + smin)) / (smax - smin)
            q = cheb_eval(coefs, mp.mpf(x)) * mp.exp(-mp.mpf(a) ** 2 / 2)
            worst = max(worst, abs(q - mp.ncdf(-mp.mpf(a))))
        print(c, deg, float(worst))
