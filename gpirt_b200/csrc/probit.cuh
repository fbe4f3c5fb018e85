// Ordinal-probit cell likelihood: log(Phi(tau_c - g) - Phi(tau_{c-1} - g) + 1e-6), the inner loop of
// ll_bar_sparse (src/log-likelihood.cpp:50-64), compute_ll_sparse (src/draw-theta.cpp:4-24) and the
// threshold ESS (src/draw_threshold.cpp:120-126).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#ifdef __CUDACC__
#include <math_constants.h>
#endif

namespace gpirt {

// ---- fast fixed-coefficient Phi and log ------------------------------------------------------------------
// The ESS kernels are bound by the fp64 pipe, not by HBM: each cell costs up to 2 Phi + 1 log per evaluation
// (CUDA libm: erfc ~100, log ~54 DFMA-equivalents, profiles/r01_microbench_fp64.txt). The likelihood only needs
// ABSOLUTE accuracy (the reference adds 1e-6 inside the log), which allows much shorter evaluations:
//   Phi(-a) = exp(-a^2/2) * P18((a-6)/(a+6)),  a in [0, 8.6]; beyond 8.6 the tail is < 1e-17 and is returned as 0
//   exp(y)  = 2^n * 2^(j/32) * E6(r),  |r| <= ln2/64                      (32-entry table)
//   log(x)  = e*ln2 + log(c_i) + L7(m/c_i - 1),  |m/c_i - 1| <= 2^-8      (128-entry table)
// Coefficients: tools/gen_probit_coeffs.py (mpmath, 60 digits). Measured against mpmath in tests/: Phi absolute
// error <= 4e-16 (R::pnorm / Cody in the reference: ~1e-16), log absolute error <= 3e-16.
#include "probit_coeffs.inc"

#ifdef __CUDACC__
#define GPIRT_TAB static __device__ const
#else
#define GPIRT_TAB static const
#endif
GPIRT_TAB double g_exp2_tab[32] = {EXP2_TAB32};
struct LogNode { double inv, logc; };
GPIRT_TAB LogNode g_log_tab[128] = {LOG_TAB128};
#ifndef __CUDA_ARCH__
static const double h_exp2_tab[32] = {EXP2_TAB32};
static const LogNode h_log_tab[128] = {LOG_TAB128};
#endif

// exp(y) for y in [-37, 0]
__host__ __device__ __forceinline__ double exp_neg(double y) {
    const double t = y * 46.16624130844683;  // 32/ln2
#ifdef __CUDA_ARCH__
    const int k = __double2int_rn(t);
#else
    const int k = (int)nearbyint(t);
#endif
    const double kd = (double)k;
    double r = fma(kd, -0.02166084938653512, y);     // ln2/32 hi (low 21 bits zero: kd*hi is exact)
    r = fma(kd, -5.9631716539705866e-12, r);  // ln2/32 lo
    double p = 1.0 / 720.0;
    p = fma(p, r, 1.0 / 120.0);
    p = fma(p, r, 1.0 / 24.0);
    p = fma(p, r, 1.0 / 6.0);
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    const int j = k & 31, n = k >> 5;  // k = 32 n + j
#ifdef __CUDA_ARCH__
    const double v = g_exp2_tab[j] * p;
    return __hiloint2double(__double2hiint(v) + (n << 20), __double2loint(v));  // * 2^n, result stays normal (n >= -54)
#else
    return ldexp(h_exp2_tab[j] * p, n);
#endif
}

// Phi(x), lower tail (R::pnorm(x, 0, 1, 1, 0) in the reference). Phi(-inf) = 0, Phi(+inf) = 1 exactly.
__host__ __device__ __forceinline__ double phi(double x) {
    const double a = fabs(x);
    if (a != a) return a;
    double q = 0.0;
    if (a <= PHI_A) {
        const double s = (a - PHI_C) / (a + PHI_C);
        const double t = fma(PHI_ALPHA, s, PHI_BETA);
#define X(v) v
        const double c[19] = {PHI_POLY(X)};
#undef X
        double p = c[18];
#pragma unroll
        for (int k = 17; k >= 0; --k) p = fma(p, t, c[k]);
        q = p * exp_neg(-0.5 * a * a);
    }
    return x < 0.0 ? q : 1.0 - q;
}

// ---- Phi from a table of grid values (round 2) --------------------------------------------------------------------------------
// Phi at full double accuracy is what bounds the log-likelihood pass: phi() above costs ~52 fp64-pipe instructions (reciprocal, degree-18
// polynomial, exp). phi_tab() takes a Taylor step of at most 1/256 from the nearest point a0 = k/128 of a table holding Q(a0) = Phi(-a0) and the density
// phi(a0): Q(a0 + d) = Q(a0) - phi(a0) * d * (1 + d (c2 + d (c3 + d (c4 + d c5)))), c_k = (-1)^(k-1) He_(k-1)(a0) / k! with the Hermite polynomials formed
// from a0 on the fly. ~27 instructions and ONE 16-byte table load; the remainder term is below 2e-18. Measured against mpmath at 50 digits over
// [-9.5, 9.5] including every half-way point between grid values: absolute error <= 1.2e-16 (phi(): 4e-16; R::pnorm in the reference: ~1e-16).
// Table: 1103 x 16 bytes (tools/gen_probit_coeffs.py -> phi_table.inc); kernels stage it in shared memory once per CTA.
#include "phi_table.inc"
struct PhiNode { double q, d; };
constexpr int PHI_TAB_N = PHI_TAB_K + 1;
GPIRT_TAB PhiNode g_phi_tab[PHI_TAB_N] = {PHI_TAB_ROWS};
#ifndef __CUDA_ARCH__
static const PhiNode h_phi_tab[PHI_TAB_N] = {PHI_TAB_ROWS};
#endif

__host__ __device__ __forceinline__ double phi_tab(double x, const PhiNode* __restrict__ tab) {
    const double a = fabs(x);
    if (a != a) return a;
    double q = 0.0;  // beyond the table (a > 8.61) the tail is < 1e-17 and is returned as 0, as phi() does
    const double t = a * PHI_TAB_H;
    if (t <= (double)PHI_TAB_K + 0.5) {
#ifdef __CUDA_ARCH__
        const int k = __double2int_rn(t);
#else
        const int k = (int)nearbyint(t);
#endif
        const double a0 = (double)k * (1.0 / PHI_TAB_H);  // exact: a multiple of 2^-7
        const double d = a - a0;                           // |d| <= 1/256
        const PhiNode nd = tab[k];
        const double a2 = a0 * a0;
        const double c2 = -0.5 * a0;
        const double c3 = fma(a2, 1.0 / 6.0, -1.0 / 6.0);
        const double c4 = -a0 * fma(a2, 1.0 / 24.0, -0.125);
        const double c5 = fma(a2 * (a2 - 6.0), 1.0 / 120.0, 0.025);
        double sser = fma(d, c5, c4);
        sser = fma(d, sser, c3);
        sser = fma(d, sser, c2);
        sser = fma(d, sser, 1.0);
        q = fma(-nd.d, sser * d, nd.q);
    }
    return x < 0.0 ? q : 1.0 - q;
}

// log(x) for positive normal x (the likelihood's argument lies in [1e-6, 1 + 1e-6]); `tab`: the 128-node table (g_log_tab, or a shared-memory copy)
__host__ __device__ __forceinline__ double log_pos_t(double x, const LogNode* __restrict__ tab) {
    // the exponent-bit arithmetic below reads NaN as ~ +-710: a NaN state must reach the ESS test as NaN (never accepted -> stall flag), as std::log
    // gives it in the reference (log(0) = -inf, log(negative) = NaN)
    if (!(x > 0.0)) return x == 0.0 ? -HUGE_VAL : x * HUGE_VAL * 0.0;
#ifdef __CUDA_ARCH__
    const long long bits = __double_as_longlong(x);
#else
    long long bits;
    memcpy(&bits, &x, 8);
#endif
    const int e = (int)(bits >> 52) - 1023;
    const int i = (int)(bits >> 45) & 127;
    const long long mb = (bits & 0x000FFFFFFFFFFFFFLL) | 0x3FF0000000000000LL;
#ifdef __CUDA_ARCH__
    const double m = __longlong_as_double(mb);
#else
    double m;
    memcpy(&m, &mb, 8);
#endif
    const LogNode nd = tab[i];
    const double r = fma(m, nd.inv, -1.0);
    double p = 1.0 / 7.0;
    p = fma(p, r, -1.0 / 6.0);
    p = fma(p, r, 0.2);
    p = fma(p, r, -0.25);
    p = fma(p, r, 1.0 / 3.0);
    p = fma(p, r, -0.5);
    p = fma(p * r, r, r);  // r + r^2 * (...)
    const double ed = (double)e;
    const double lo = fma(ed, 4.7493250390316726e-07, p);    // ln2 lo
    const double hi = fma(ed, 0.6931467056274414, nd.logc);  // ln2 hi (low 32 bits zero: ed*hi is exact)
    return hi + lo;
}

__host__ __device__ __forceinline__ double log_pos(double x) {
#ifdef __CUDA_ARCH__
    return log_pos_t(x, g_log_tab);
#else
    return log_pos_t(x, h_log_tab);
#endif
}

// log(Phi_hi - Phi_lo + 1e-6) with the reference's operation order (src/log-likelihood.cpp:62-63)
__host__ __device__ __forceinline__ double cell_from_probs(double p_lo, double p_hi) {
#ifdef __CUDA_ARCH__
    return log_pos(__dadd_rn(__dsub_rn(p_hi, p_lo), 1e-6));
#else
    return log_pos((p_hi - p_lo) + 1e-6);
#endif
}

// c in 1..C; tau_lo = tau[c-1], tau_hi = tau[c]; edge categories skip the half that is exactly 0 or 1.
__host__ __device__ __forceinline__ double cell_ll(double g, double tau_lo, double tau_hi, bool first, bool last) {
    double p_hi = last ? 1.0 : phi(tau_hi - g);
    double p_lo = first ? 0.0 : phi(tau_lo - g);
    return cell_from_probs(p_lo, p_hi);
}

// the same cell term with Phi from the table (k_item_ll, k_theta_table)
__host__ __device__ __forceinline__ double cell_ll_tab(double g, double tau_lo, double tau_hi, bool first, bool last, const PhiNode* __restrict__ tab) {
    double p_hi = last ? 1.0 : phi_tab(tau_hi - g, tab);
    double p_lo = first ? 0.0 : phi_tab(tau_lo - g, tab);
    return cell_from_probs(p_lo, p_hi);
}
// Branch-free form for a uniform instruction stream: both tails are always evaluated. The open ends of an item's thresholds are -inf / +inf
// (src/log-likelihood.cpp:68-70), for which phi_tab returns exactly 0 / 1 — the values the branches above substitute — so the result is the same bits.
__host__ __device__ __forceinline__ double cell_ll_tab2(double g, double tau_lo, double tau_hi, const PhiNode* __restrict__ tab, const LogNode* __restrict__ ltab) {
    const double p_hi = phi_tab(tau_hi - g, tab);
    const double p_lo = phi_tab(tau_lo - g, tab);
#ifdef __CUDA_ARCH__
    return log_pos_t(__dadd_rn(__dsub_rn(p_hi, p_lo), 1e-6), ltab);
#else
    return log_pos_t((p_hi - p_lo) + 1e-6, ltab);
#endif
}

#ifdef __CUDACC__
// stage the Phi table in shared memory (every thread of the CTA calls this; a barrier follows at the caller)
__device__ __forceinline__ void load_phi_table(PhiNode* s_tab) {
    for (int q = threadIdx.x + blockDim.x * threadIdx.y; q < PHI_TAB_N; q += blockDim.x * blockDim.y) s_tab[q] = g_phi_tab[q];
}

// ---- fp32 screening of ESS proposals ---------------------------------------------------------------------------
// An ESS proposal is accepted iff ll' > log_y. Most proposals miss log_y by far more than fp32 can blur, so the kernels first
// evaluate ll' in fp32 TOGETHER WITH a rigorous bound E on |ll'_fp32 - ll'_fp64|, and only fall back to the fp64 evaluation when
// |ll'_fp32 - log_y| <= E. The decision — hence every accepted state, every proposal count and every uniform consumed — is
// exactly the fp64 decision (tests: screening on vs off is bit-identical); the fp32 value itself is never stored.
// Per cell, with a = tau_lo - g, b = tau_hi - g (formed in fp64, then rounded), u = |a|, v = |b| and Q the upper tail:
//   Phi(b) - Phi(a) =  Q(u) - Q(v)        if a >= 0      (both in the upper tail: no cancellation against 1)
//                      Q(v) - Q(u)        if b <= 0
//                      1 - Q(u) - Q(v)    otherwise
//   Q(x) = 2^(-x^2 log2(e)/2) * P7((x-4)/(x+4)), x clamped to 9.25 (Q < 1.2e-20 beyond), relative error rho(x) <= 4e-6 + 4e-7 x^2
//   (polynomial 4e-7, fp32 Horner ~1e-6, ex2.approx 2^-22, rounding of x and of the exponent: 1.8e-7 x^2; 2x margin)
//   dp = rho(u) Q(u) + rho(v) Q(v) + 4 ulp32 (p + [straddle])        bound on the error of p = Phi(b) - Phi(a) + 1e-6
//   e  = dp / (p - dp) + 2e-6 + 5e-7 |t|                              bound on the error of t = log p (lg2.approx: 2^-22)
// tools/emul_screen32.py reproduces this arithmetic in numpy over 2e6 random cells: observed error / bound <= 0.29.
__device__ __forceinline__ float f_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float f_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float f_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ float q_tail32(float x) {  // x already clamped to [0, 9.25]
    const float s = (x - 4.0f) * f_rcp(x + 4.0f);
    float p = 0.000785236840738696f;
    p = fmaf(p, s, 0.0005041266204986191f);
    p = fmaf(p, s, -0.008529436410853592f);
    p = fmaf(p, s, 0.03085854142686483f);
    p = fmaf(p, s, -0.07170505823364434f);
    p = fmaf(p, s, 0.12437966746436852f);
    p = fmaf(p, s, -0.1703977371436323f);
    p = fmaf(p, s, 0.09441063861780297f);
    return p * f_ex2(x * x * -0.72134752f);
}

// t ~ log(Phi(tau_hi - g) - Phi(tau_lo - g) + 1e-6) in fp32 and e >= |t - (the fp64 cell term)|. Infinite thresholds (edge
// categories) need no special case: their tail clamps to Q(9.25) ~ 1e-20. A NaN anywhere gives e = NaN, which fails both
// screening comparisons and sends the proposal to the fp64 path.
// the cell term from the two rounded arguments a = tau_lo - g, b = tau_hi - g and their tails qu = Q(min(|a|, 9.25)), qv = Q(min(|b|, 9.25))
__device__ __forceinline__ void cell_ll32_core(float a, float b, float u, float v, float qu, float qv, float& t, float& e) {
    const bool sa = a >= 0.0f, sb = b <= 0.0f;
    const float straddle = (sa || sb) ? 0.0f : 1.0f;
    const float d = sa ? qu - qv : (sb ? qv - qu : (1.0f - qu) - qv);
    const float p = d + 1e-6f;
    t = f_lg2(p) * 0.6931472f;
    const float dp = fmaf(fmaf(4e-7f, u * u, 4e-6f), qu, fmaf(4e-7f, v * v, 4e-6f) * qv) + 2.4e-7f * (p + straddle);
    const float den = p - dp;
    e = (den > dp ? dp * f_rcp(den) * 1.0001f : 14.0f) + fmaf(5e-7f, fabsf(t), 2e-6f);
    // fminf(|NaN|, 9.25) is 9.25: a NaN argument (NaN in f / beta / a threshold, or inf - inf) would otherwise pass as a deep-tail cell with a small bound.
    // A NaN bound fails both screening comparisons, so the proposal goes to the fp64 pass, which carries the NaN into the ESS test (never accepted -> stall flag).
    if (a != a || b != b) e = CUDART_NAN_F;
}
__device__ __forceinline__ void cell_ll32(double g, double tau_lo, double tau_hi, float& t, float& e) {
    const float a = (float)__dsub_rn(tau_lo, g), b = (float)__dsub_rn(tau_hi, g);
    const float u = fminf(fabsf(a), 9.25f), v = fminf(fabsf(b), 9.25f);
    cell_ll32_core(a, b, u, v, q_tail32(u), q_tail32(v), t, e);
}

// ---- two cells per instruction stream: sm_100's packed fp32 pairs (FADD2 / FMUL2 / FFMA2: add / mul / fma .rn.f32x2) --------------------------
// The ESS item kernels are bound by instruction issue (ncu: issue slots 77 % busy, every pipe below 40 %, profiles/r02_ncu_ess.txt), and almost all
// of a cell's fp32 arithmetic is the same operation on its own two tails and on the next cell's. A pair holds lane 0 = cell 0, lane 1 = cell 1: the
// lower tails of both cells form one pair, the upper tails another. Every lane operation is the IEEE round-to-nearest operation of the scalar code
// above, in the same order, so (t, e) are the bits cell_ll32 gives (gpirt_debug_cell32 checks that) — one instruction for two cells instead of two.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) { f32x2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 bc2(float c) { return pk2(c, c); }
__device__ __forceinline__ float min_nan(float a, float b) { float r; asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }

// Q of both lanes (arguments already clamped to [0, 9.25]); xx receives the squares (the error weights reuse them)
__device__ __forceinline__ f32x2 q_tail32x2(f32x2 x, f32x2& xx) {
    const f32x2 xm = add2(x, bc2(-4.0f)), xp = add2(x, bc2(4.0f));
    float p0, p1;
    upk2(xp, p0, p1);
    const f32x2 s = mul2(xm, pk2(f_rcp(p0), f_rcp(p1)));
    f32x2 p = bc2(0.000785236840738696f);
    p = fma2(p, s, bc2(0.0005041266204986191f));
    p = fma2(p, s, bc2(-0.008529436410853592f));
    p = fma2(p, s, bc2(0.03085854142686483f));
    p = fma2(p, s, bc2(-0.07170505823364434f));
    p = fma2(p, s, bc2(0.12437966746436852f));
    p = fma2(p, s, bc2(-0.1703977371436323f));
    p = fma2(p, s, bc2(0.09441063861780297f));
    xx = mul2(x, x);
    float e0, e1;
    upk2(mul2(xx, bc2(-0.72134752f)), e0, e1);
    return mul2(p, pk2(f_ex2(e0), f_ex2(e1)));
}

// cells 0 and 1 at once; same results as two calls of cell_ll32
__device__ __forceinline__ void cell_ll32x2(double g0, double lo0, double hi0, double g1, double lo1, double hi1, float& t0, float& e0, float& t1,
                                            float& e1) {
    const float a0 = (float)__dsub_rn(lo0, g0), b0 = (float)__dsub_rn(hi0, g0);
    const float a1 = (float)__dsub_rn(lo1, g1), b1 = (float)__dsub_rn(hi1, g1);
    // min.NaN keeps a NaN argument (fminf would clamp it to 9.25 and pass it as a deep-tail cell): it then runs through both tails into t and e of its
    // cell — every form of d below uses both tails — so the cell needs no NaN test of its own: a NaN sum or bound fails both screening comparisons
    const f32x2 U = pk2(min_nan(fabsf(a0), 9.25f), min_nan(fabsf(a1), 9.25f));  // the lower thresholds' tails of both cells
    const f32x2 V = pk2(min_nan(fabsf(b0), 9.25f), min_nan(fabsf(b1), 9.25f));  // the upper thresholds'
    f32x2 UU, VV;
    const f32x2 QU = q_tail32x2(U, UU), QV = q_tail32x2(V, VV);
    // the three forms of Phi(b) - Phi(a) (cell_ll32_core), then each cell picks its own
    const f32x2 X1 = sub2(QU, QV);                    // both arguments in the upper tail:  qu - qv   (negated: both in the lower tail)
    const f32x2 X3 = sub2(sub2(bc2(1.0f), QU), QV);   // the bin straddles zero:            (1 - qu) - qv
    float x10, x11, x30, x31;
    upk2(X1, x10, x11);
    upk2(X3, x30, x31);
    const bool sa0 = a0 >= 0.0f, sb0 = b0 <= 0.0f, sa1 = a1 >= 0.0f, sb1 = b1 <= 0.0f;
    const float d0 = sa0 ? x10 : (sb0 ? -x10 : x30), d1 = sa1 ? x11 : (sb1 ? -x11 : x31);
    const f32x2 S = pk2((sa0 || sb0) ? 0.0f : 1.0f, (sa1 || sb1) ? 0.0f : 1.0f);
    const f32x2 P = add2(pk2(d0, d1), bc2(1e-6f));
    float p0, p1;
    upk2(P, p0, p1);
    t0 = f_lg2(p0) * 0.6931472f;
    t1 = f_lg2(p1) * 0.6931472f;
    // dp = (4e-7 u^2 + 4e-6) qu + (4e-7 v^2 + 4e-6) qv + 2.4e-7 (p + straddle), with cell_ll32_core's association
    const f32x2 RU = fma2(bc2(4e-7f), UU, bc2(4e-6f)), RV = fma2(bc2(4e-7f), VV, bc2(4e-6f));
    const f32x2 DP = add2(fma2(RU, QU, mul2(RV, QV)), mul2(bc2(2.4e-7f), add2(P, S)));
    const f32x2 DEN = sub2(P, DP);
    float dp0, dp1, den0, den1;
    upk2(DP, dp0, dp1);
    upk2(DEN, den0, den1);
    const float r0 = den0 > dp0 ? dp0 * f_rcp(den0) * 1.0001f : 14.0f, r1 = den1 > dp1 ? dp1 * f_rcp(den1) * 1.0001f : 14.0f;
    e0 = r0 + fmaf(5e-7f, fabsf(t0), 2e-6f);  // t is NaN for a NaN argument, and so is e
    e1 = r1 + fmaf(5e-7f, fabsf(t1), 2e-6f);
}
#endif

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
