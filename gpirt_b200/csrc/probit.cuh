// Ordinal-probit cell likelihood: log(Phi(tau_c - g) - Phi(tau_{c-1} - g) + 1e-6), the inner loop of
// ll_bar_sparse (src/log-likelihood.cpp:50-64), compute_ll_sparse (src/draw-theta.cpp:4-24) and the
// threshold ESS (src/draw_threshold.cpp:120-126).
#pragma once
#include <cstdint>
#include <math_constants.h>

namespace gpirt {

// Phi(x), lower tail. R::pnorm(x,0,1,1,0) in the reference (Cody's rational approximations, ~1e-16);
// here 0.5*erfc(-x/sqrt2) via the CUDA fp64 erfc (<= 5 ulp). Phi(-inf) = 0 and Phi(+inf) = 1 exactly.
__device__ __forceinline__ double phi(double x) {
    return 0.5 * erfc(-0.70710678118654752440 * x);
}

// c in 1..C; tau_lo = tau[c-1], tau_hi = tau[c]; edge categories skip the half that is exactly 0 or 1.
__device__ __forceinline__ double cell_ll(double g, double tau_lo, double tau_hi, bool first, bool last) {
    double p_hi = last ? 1.0 : phi(tau_hi - g);
    double p_lo = first ? 0.0 : phi(tau_lo - g);
    return log(p_hi - p_lo + 1e-6);
}

// grid index of theta on theta* = -5:0.01:5 (src/draw-theta.cpp:10-11, src/gpirtMCMC.cpp:291-296):
// round-half-away((theta + 5.0) / 0.01), clamped. Explicit IEEE add/div so no contraction can change it.
__host__ __device__ __forceinline__ int grid_index(double theta) {
#ifdef __CUDA_ARCH__
    double q = __ddiv_rn(__dadd_rn(theta, 5.0), 0.01);
#else
    double q = (theta + 5.0) / 0.01;
#endif
    int idx = (int)round(q);
    return idx < 0 ? 0 : (idx > 1000 ? 1000 : idx);
}
// theta*[idx] as arma::regspace(-5.0, 0.01, 5.0) produces it: start + idx*delta
__host__ __device__ __forceinline__ double grid_value(int idx) {
#ifdef __CUDA_ARCH__
    return __dadd_rn(-5.0, __dmul_rn((double)idx, 0.01));
#else
    return -5.0 + (double)idx * 0.01;
#endif
}

// mu = X beta for one cell: [1, t, t^2] . (b0, b1, b2), accumulated left to right without contraction
__device__ __forceinline__ double mu_of(double t, double b0, double b1, double b2) {
    return __dadd_rn(__dadd_rn(b0, __dmul_rn(t, b1)), __dmul_rn(__dmul_rn(t, t), b2));
}

}  // namespace gpirt
