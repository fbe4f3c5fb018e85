"""Numpy float32 emulation of the fp32 screening arithmetic of csrc/probit.cuh (cell_ll32): fits the degree-7 polynomial of the
Mills-type factor Q(x) exp(x^2/2) in s = (x-4)/(x+4) and checks observed error / claimed bound over 2e6 random cells."""
import numpy as np, mpmath as mp
from scipy.special import ndtr, log_ndtr
f32=np.float32
mp.mp.dps=40
K=4.0; A=9.25; deg=7
def Rm(u): u=mp.mpf(u); return mp.ncdf(-u)*mp.exp(u*u/2)
s0=-1.0; s1=(A-K)/(A+K); N=deg+1
xs=np.cos(np.pi*(np.arange(N)+0.5)/N); ss=0.5*(s1+s0)+0.5*(s1-s0)*xs; us=K*(1+ss)/(1-ss)
ys=np.array([float(Rm(u)) for u in us])
c=np.polynomial.chebyshev.chebfit(xs,ys,deg); p=np.polynomial.chebyshev.cheb2poly(c)
# convert poly in x to poly in s: x = (s - mid)/half
mid=0.5*(s1+s0); half=0.5*(s1-s0)
ps=np.polynomial.polynomial.Polynomial(p)(np.polynomial.polynomial.Polynomial([-mid/half,1/half])).coef
print("coef s:",[float(v) for v in ps])
cf=[f32(v) for v in ps]
def Q32(u):  # u float32 array in [0,A]
    u=np.minimum(u,f32(A))
    r=(f32(1)/(u+f32(K))).astype(f32)
    s=((u-f32(K))*r).astype(f32)
    acc=np.full_like(u,cf[deg])
    for k in range(deg-1,-1,-1): acc=(acc*s+cf[k]).astype(f32)
    arg=((u*u).astype(f32)*f32(-0.72134752)).astype(f32)
    e=np.exp2(arg.astype(np.float64)).astype(f32)
    return (acc*e).astype(f32)
rng=np.random.default_rng(0)
n=2_000_000
g=rng.normal(0,3,n); 
tl=rng.normal(0,2,n); w=np.exp(rng.normal(-0.5,1.5,n)); th=tl+w
kind=rng.integers(0,3,n)  # 0 interior,1 first,2 last
tl=np.where(kind==1,-np.inf,tl); th=np.where(kind==2,np.inf,th)
a=tl-g; b=th-g
# fp64 truth
d64=np.where(a>=0, ndtr(-a)-ndtr(-b), ndtr(b)-ndtr(a))
t64=np.log(d64+1e-6)
a32=a.astype(f32); b32=b.astype(f32)
u=np.abs(a32); v=np.abs(b32)
Qu=Q32(u); Qv=Q32(v)
sa=a32>=0; sb=b32<=0
strad=(~sa)&(~sb)
d=np.where(sa,(Qu-Qv).astype(f32),np.where(sb,(Qv-Qu).astype(f32),((f32(1)-Qu).astype(f32)-Qv).astype(f32)))
p32=(d+f32(1e-6)).astype(f32)
t32=(np.log2(p32.astype(np.float64))).astype(f32)*f32(0.6931472)
t32=t32.astype(f32)
uc=np.minimum(u,f32(A)); vc=np.minimum(v,f32(A))
rho_u=(f32(4e-6)+f32(4e-7)*uc*uc); rho_v=(f32(4e-6)+f32(4e-7)*vc*vc)
dp=rho_u*Qu+rho_v*Qv+f32(4*2.0**-24)*(p32+strad.astype(f32))
den=p32-dp
e=np.where(den>dp, dp/den, 14.0)+4e-6
err=np.abs(t32.astype(np.float64)-t64)
ratio=err/e
print("max err",err.max(),"max ratio err/e",ratio.max(), "mean e",e.mean(), "frac capped",(den<=dp).mean())
i=np.argmax(ratio); print(a[i],b[i],t64[i],t32[i],e[i],Qu[i],Qv[i],p32[i],dp[i])
# typical E for n=2000 sum
print("typical E per 2000 cells:", e[:2000].sum(), e[2000:4000].sum())
