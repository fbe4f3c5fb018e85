// TEST INFRASTRUCTURE — CPU oracle for the gpirt per-sweep hot path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// use anything under oracle/. The product (gpirt_b200/) never includes, links or calls this.
//
// Numerical leaves restated from the reference (citations relative to /root/reference):
//   K, K_time                  src/covariance-function.cpp:3-44
//   chol / double_solve        src/cholesky-cache.cpp:4-24 (Armadillo -> LAPACK dpotrf/dtrtrs there;
//                              plain unblocked loops here, plus a long-double variant for kappa-aware checks)
//   ll_bar_sparse              src/log-likelihood.cpp:50-64
//   delta_to_threshold etc.    src/log-likelihood.cpp:66-88
//   R::pnorm                   third-party (Rmath nmath/pnorm.c, unpinned: DESCRIPTION:25-26). Restated
//                              here from W. J. Cody (1969) "Rational Chebyshev approximations for the
//                              error function", the algorithm R documents for pnorm; pinned in
//                              tests/test_oracle_math.py against mpmath at 40 digits.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <type_traits>
#include <vector>

#include "orc_blas.h"

namespace orc {

// ---------------------------------------------------------------------------------------------
// Phi(x): lower tail, non-log. Cody's three-region rational approximation as used by Rmath.
inline double pnorm(double x) {
    static const double a[5] = {2.2352520354606839287, 161.02823106855587881, 1067.6894854603709582,
                                18154.981253343561249, 0.065682337918207449113};
    static const double b[4] = {47.20258190468824187, 976.09855173777669322, 10260.932208618978205,
                                45507.789335026729956};
    static const double c[9] = {0.39894151208813466764, 8.8831497943883759412, 93.506656132177855979,
                                597.27027639480026226,  2494.5375852903726711, 6848.1904505362823326,
                                11602.651437647350124,  9842.7148383839780218, 1.0765576773720192317e-8};
    static const double d[8] = {22.266688044328115691, 235.38790178262499861, 1519.377599407554805,
                                6485.558298266760755,  18615.571640885098091, 34900.952721145977266,
                                38912.003286093271411, 19685.429676859990727};
    static const double p[6] = {0.21589853405795699,     0.1274011611602473639, 0.022235277870649807,
                                0.001421619193227893466, 2.9112874951168792e-5, 0.02307344176494017303};
    static const double q[5] = {1.28426009614491121, 0.468238212480865118, 0.0659881378689285515,
                                0.00378239633202758244, 7.29751555083966205e-5};
    const double SQRT32 = 5.656854249492380195206754896838;
    const double INV_SQRT_2PI = 0.398942280401432677939946059934;
    if (std::isnan(x)) return x;
    const double eps = std::numeric_limits<double>::epsilon() * 0.5;
    double y = std::fabs(x), cum, ccum, xnum, xden, temp, xsq, del;
    if (y <= 0.67448975) {
        if (y > eps) {
            xsq = x * x;
            xnum = a[4] * xsq;
            xden = xsq;
            for (int i = 0; i < 3; ++i) { xnum = (xnum + a[i]) * xsq; xden = (xden + b[i]) * xsq; }
        } else {
            xnum = xden = 0.0;
        }
        temp = x * (xnum + a[3]) / (xden + b[3]);
        return 0.5 + temp;
    } else if (y <= SQRT32) {
        xnum = c[8] * y;
        xden = y;
        for (int i = 0; i < 7; ++i) { xnum = (xnum + c[i]) * y; xden = (xden + d[i]) * y; }
        temp = (xnum + c[7]) / (xden + d[7]);
        xsq = std::trunc(y * 16) / 16;
        del = (y - xsq) * (y + xsq);
        cum = std::exp(-xsq * xsq * 0.5) * std::exp(-del * 0.5) * temp;
        ccum = 1.0 - cum;
        return x > 0 ? ccum : cum;
    } else if (x > -37.5193 && x < 8.2924) {
        xsq = 1.0 / (x * x);
        xnum = p[5] * xsq;
        xden = xsq;
        for (int i = 0; i < 4; ++i) { xnum = (xnum + p[i]) * xsq; xden = (xden + q[i]) * xsq; }
        temp = xsq * (xnum + p[4]) / (xden + q[4]);
        temp = (INV_SQRT_2PI - temp) / y;
        xsq = std::trunc(x * 16) / 16;
        del = (x - xsq) * (x + xsq);
        cum = std::exp(-xsq * xsq * 0.5) * std::exp(-del * 0.5) * temp;
        ccum = 1.0 - cum;
        return x > 0 ? ccum : cum;
    }
    return x > 0 ? 1.0 : 0.0;
}

// ---------------------------------------------------------------------------------------------
// K(x1,x2,sds): out is n1 x n2 column-major. Operation order follows src/covariance-function.cpp:10-14.
inline void K(const double* x1, size_t n1, const double* x2, size_t n2, const double* sds, double* out) {
    for (size_t j = 0; j < n2; ++j) {
        for (size_t i = 0; i < n1; ++i) {
            double diff = x1[i] - x2[j];
            double r = std::exp(-0.5 * diff * diff);
            r += x1[i] * std::pow(sds[1], 2) * x2[j];
            r += std::pow(sds[0], 2);
            r += std::pow(x1[i] * sds[2] * x2[j], 2);
            out[i + j * n1] = r;
        }
    }
}

// K_time: Matern-5/2 (kernel==0) or RBF (kernel==1) over session indices (src/covariance-function.cpp:21-44).
inline void K_time(const double* x1, size_t n1, const double* x2, size_t n2, double os, double ls,
                   const double* tsds /*2*/, int kernel, double* out) {
    for (size_t j = 0; j < n2; ++j) {
        for (size_t i = 0; i < n1; ++i) {
            double diff = std::fabs(x1[i] - x2[j]);
            double r = 0.0;
            if (kernel == 0) {
                r = os * os * (1 + std::sqrt(5) * diff / ls + 5 * diff * diff / ls / ls / 3);
                r = r * std::exp(-std::sqrt(5) * diff / ls);
                r += x1[i] * std::pow(tsds[1], 2) * x2[j];
                r += std::pow(tsds[0], 2);
            } else if (kernel == 1) {
                r = os * os * std::exp(-diff * diff / ls / ls);
                r += x1[i] * std::pow(tsds[1], 2) * x2[j];
                r += std::pow(tsds[0], 2);
            }
            out[i + j * n1] = r;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Lower Cholesky, column-major, in place (upper triangle zeroed as arma::chol(.,"lower") returns).
// Returns 0 or the 1-based index of the failing leading minor. T = double or long double.
template <typename T>
inline int chol_lower(T* A, size_t n) {
    for (size_t j = 0; j < n; ++j) {
        T* cj = A + j * n;
        // left-looking column update: A[j:,j] -= L[j:, :j] * L[j, :j]^T
        for (size_t k = 0; k < j; ++k) {
            const T* ck = A + k * n;
            T ljk = ck[j];
            if (ljk == T(0)) continue;
            for (size_t i = j; i < n; ++i) cj[i] -= ck[i] * ljk;
        }
        T d = cj[j];
        if (!(d > T(0))) return int(j + 1);
        d = std::sqrt(d);
        cj[j] = d;
        for (size_t i = j + 1; i < n; ++i) cj[i] /= d;
        for (size_t i = 0; i < j; ++i) cj[i] = T(0);
    }
    return 0;
}

// Blocked right-looking variant of the same factorisation (64-wide panels, OpenMP over the trailing
// tiles): what the timed CPU baseline uses so that the O(n^3) part is cache-blocked and threaded, as
// the LAPACK dpotrf behind arma::chol would be. Same arithmetic up to summation order.
inline int chol_lower_blocked(double* A, size_t n) {
    if (blas().on) {  // timed baseline only (orc_blas.h): LAPACK dpotrf, as arma::chol(S, "lower") calls it; upper triangle zeroed like Armadillo does
        int nn = int(n), info = 0;
        blas().dpotrf("L", &nn, A, &nn, &info);
        for (size_t j = 0; j < n; ++j)
            for (size_t i = 0; i < j; ++i) A[i + j * n] = 0.0;
        return info;
    }
    const size_t NB = 64;
    if (n <= 2 * NB) return chol_lower(A, n);
    for (size_t k0 = 0; k0 < n; k0 += NB) {
        const size_t kb = std::min(NB, n - k0);
        // diagonal block (unblocked, in place inside A)
        for (size_t j = k0; j < k0 + kb; ++j) {
            double* cj = A + j * n;
            for (size_t k = k0; k < j; ++k) {
                const double* ck = A + k * n;
                const double ljk = ck[j];
                for (size_t i = j; i < k0 + kb; ++i) cj[i] -= ck[i] * ljk;
            }
            double d = cj[j];
            if (!(d > 0.0)) return int(j + 1);
            d = std::sqrt(d);
            cj[j] = d;
            for (size_t i = j + 1; i < k0 + kb; ++i) cj[i] /= d;
        }
        const size_t r0 = k0 + kb;
        if (r0 >= n) break;
        // panel: rows r0.. of columns k0..k0+kb solve X L11^T = A21 (row chunks in parallel)
#pragma omp parallel for schedule(static)
        for (size_t rb = r0; rb < n; rb += NB) {
            const size_t re = std::min(rb + NB, n);
            for (size_t j = k0; j < k0 + kb; ++j) {
                double* cj = A + j * n;
                for (size_t k = k0; k < j; ++k) {
                    const double* ck = A + k * n;
                    const double ljk = ck[j];
                    for (size_t i = rb; i < re; ++i) cj[i] -= ck[i] * ljk;
                }
                const double d = cj[j];
                for (size_t i = rb; i < re; ++i) cj[i] /= d;
            }
        }
        // trailing update A22 -= L21 L21^T, lower tiles only
        const size_t nt = (n - r0 + NB - 1) / NB;
#pragma omp parallel for schedule(dynamic)
        for (size_t t = 0; t < nt * nt; ++t) {
            const size_t bi = t % nt, bj = t / nt;
            if (bi < bj) continue;
            const size_t i0 = r0 + bi * NB, i1 = std::min(i0 + NB, n), j0 = r0 + bj * NB, j1 = std::min(j0 + NB, n);
            for (size_t j = j0; j < j1; ++j) {
                double* cj = A + j * n;
                for (size_t k = k0; k < k0 + kb; ++k) {
                    const double* ck = A + k * n;
                    const double ljk = ck[j];
                    for (size_t i = std::max(i0, j); i < i1; ++i) cj[i] -= ck[i] * ljk;
                }
            }
        }
    }
    for (size_t j = 0; j < n; ++j)
        for (size_t i = 0; i < j; ++i) A[i + j * n] = 0.0;
    return 0;
}

// forward substitution L x = b (in place), back substitution L^T x = b (in place)
template <typename T>
inline void trsv_lower(const T* L, size_t n, T* x) {
    if constexpr (std::is_same<T, double>::value) {
        if (blas().on) { blas().dtrsv(102, 122, 111, 131, int(n), L, int(n), x, 1); return; }
    }
    // column-oriented (axpy form): streams L column by column
    for (size_t j = 0; j < n; ++j) {
        T xj = x[j] / L[j + j * n];
        x[j] = xj;
        const T* cj = L + j * n;
        for (size_t i = j + 1; i < n; ++i) x[i] -= cj[i] * xj;
    }
}
template <typename T>
inline void trsv_lower_t(const T* L, size_t n, T* x) {
    if constexpr (std::is_same<T, double>::value) {
        if (blas().on) { blas().dtrsv(102, 122, 112, 131, int(n), L, int(n), x, 1); return; }
    }
    for (size_t jj = n; jj-- > 0;) {
        const T* cj = L + jj * n;
        T s = x[jj];
        for (size_t i = jj + 1; i < n; ++i) s -= cj[i] * x[i];
        x[jj] = s / cj[jj];
    }
}
// double_solve (src/cholesky-cache.cpp:4-9): L^-T (L^-1 x)
inline void double_solve(const double* L, size_t n, double* x) {
    trsv_lower(L, n, x);
    trsv_lower_t(L, n, x);
}

// dense y = A x, A is r x c column-major (src/mvnormal.h:23 uses a dense product, ignoring triangularity)
inline void gemv(const double* A, size_t r, size_t c, const double* x, double* y) {
    if (blas().on) { blas().dgemv(102, 111, int(r), int(c), 1.0, A, int(r), x, 1, 0.0, y, 1); return; }
    for (size_t i = 0; i < r; ++i) y[i] = 0.0;
    for (size_t j = 0; j < c; ++j) {
        const double xj = x[j];
        const double* cj = A + j * r;
        for (size_t i = 0; i < r; ++i) y[i] += cj[i] * xj;
    }
}

// ---------------------------------------------------------------------------------------------
// threshold transforms (src/log-likelihood.cpp:66-88). deltas has C-1 entries, thresholds C+1.
inline void delta_to_threshold(const double* deltas, size_t C, double* thr) {
    for (size_t i = 0; i <= C; ++i) thr[i] = 0.0;
    thr[0] = -std::numeric_limits<double>::infinity();
    thr[1] = deltas[0];
    thr[C] = std::numeric_limits<double>::infinity();
    for (size_t i = 1; i + 1 < C; ++i) thr[i + 1] = thr[i] + std::exp(deltas[i]);
}
inline void threshold_to_delta(const double* thr, size_t C, double* deltas) {
    for (size_t i = 0; i + 1 < C; ++i) deltas[i] = 0.0;
    deltas[0] = thr[1];
    for (size_t i = 1; i + 1 < C; ++i) deltas[i] = std::log(thr[i + 1] - thr[i]);
}

// one ordinal-probit cell term: log(Phi(thr[c]-g) - Phi(thr[c-1]-g) + 1e-6)
inline double cell_ll(double g, int c, const double* thr, size_t thr_stride = 1) {
    double z1 = thr[(c - 1) * thr_stride] - g;
    double z2 = thr[c * thr_stride] - g;
    return std::log(pnorm(z2) - pnorm(z1) + 1e-6);
}

// ll_bar_sparse (src/log-likelihood.cpp:50-64). y: doubles, NaN = missing; obs = indices of finite y.
inline double ll_bar_sparse(const double* f, const double* y, const double* mu, const double* thr,
                            const uint64_t* obs, size_t n_obs) {
    double result = 0.0;
    for (size_t k = 0; k < n_obs; ++k) {
        size_t i = obs[k];
        double g = f[i] + mu[i];
        result += cell_ll(g, int(y[i]), thr);
    }
    return result;
}

}  // namespace orc
