import mpmath as mp, numpy as np
mp.mp.dps = 50
A = 8.6
def R(a):
    a = mp.mpf(a); return mp.ncdf(-a) * mp.exp(a * a / 2)
def fit(c, deg, npts=300):
    c = mp.mpf(c)
    smin, smax = mp.mpf(-1), (A - c) / (A + c)
    ys, xs = [], []
    for kk in range(npts):
        x = mp.cos(mp.pi * (kk + mp.mpf(1)/2) / npts)
        s = (smax - smin) / 2 * x + (smax + smin) / 2
        a = c * (1 + s) / (1 - s)
        ys.append(R(a)); xs.append(x)
    coefs = []
    for j in range(deg + 1):
        acc = mp.mpf(0)
        for kk in range(npts):
            acc += ys[kk] * mp.cos(mp.pi * j * (kk + mp.mpf(1)/2) / npts)
        coefs.append(acc * 2 / npts)
    coefs[0] /= 2
    return coefs, smin, smax
def cheb_eval(coefs, x):
    b1 = b2 = mp.mpf(0)
    for cj in reversed(coefs[1:]):
        b1, b2 = 2 * x * b1 - b2 + cj, b1
    return x * b1 - b2 + coefs[0]
for c in (4.0, 5.0, 6.0, 8.0):
    for deg in (12, 14, 16, 18):
        coefs, smin, smax = fit(c, deg)
        worst = 0
        for a in np.linspace(0, A, 800):
            a = mp.mpf(float(a)); s = (a - c) / (a + c)
            x = (2 * s - (smax + smin)) / (smax - smin)
            q = cheb_eval(coefs, x) * mp.exp(-a * a / 2)
            worst = max(worst, abs(q - mp.ncdf(-a)))
        print(c, deg, float(worst))
