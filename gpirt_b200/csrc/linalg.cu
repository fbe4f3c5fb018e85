// Dense fp64 building blocks of the sweep: covariance construction (src/covariance-function.cpp:3-44),
// blocked lower Cholesky (arma::chol(S,"lower"), src/cholesky-cache.cpp:23) and the triangular solves
// behind double_solve / solve(trimatl(L), kstar) (src/cholesky-cache.cpp:4-9, src/draw-fstar.cpp:30).
#include "linalg.cuh"

#include <cstdio>

namespace gpirt {

long long g_kernel_launches = 0;

// ------------------------------------------------------------------------------------------------
// GEMM dispatch
template <int BM, int BN, int BK, int WM, int WN, int STAGES, bool TA, bool TB>
static cudaError_t launch_gemm(cudaStream_t st, const GemmArgs& g) {
    using Cfg = GemmCfg<BM, BN, BK, WM, WN, STAGES, TA, TB>;
    auto kern = gemm_f64_kernel<BM, BN, BK, WM, WN, STAGES, TA, TB>;
    static bool configured = false;  // per instantiation
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    dim3 grid((g.M + BM - 1) / BM, (g.N + BN - 1) / BN);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(g);
    ++g_kernel_launches;
    return cudaGetLastError();
}

template <bool TA, bool TB>
static cudaError_t gemm_pick(cudaStream_t st, const GemmArgs& g) {
    // big tiles only when they still fill the 148 SMs; otherwise 64x64 tiles (4x the CTAs)
    long tiles128 = (long)((g.M + 127) / 128) * ((g.N + 127) / 128);
    if (g.lower_only) tiles128 = tiles128 / 2 + 1;
    if (tiles128 >= 2 * 148 && g.N >= 128 && g.M >= 128) return launch_gemm<128, 128, 16, 64, 32, 3, TA, TB>(st, g);
    return launch_gemm<64, 64, 16, 32, 32, 3, TA, TB>(st, g);
}

cudaError_t gemm_f64(cudaStream_t st, bool ta, bool tb, const GemmArgs& g) {
    if (g.M <= 0 || g.N <= 0) return cudaSuccess;
    if (ta) return tb ? gemm_pick<true, true>(st, g) : gemm_pick<true, false>(st, g);
    return tb ? gemm_pick<false, true>(st, g) : gemm_pick<false, false>(st, g);
}

// ------------------------------------------------------------------------------------------------
// K(x1, x2, sds): out[i + j*ld] = exp(-0.5 d^2) + x1_i*s1^2*x2_j + s0^2 + (x1_i*s2*x2_j)^2, same operation
// order as src/covariance-function.cpp:10-14, explicit IEEE mul/add (no contraction); diag_add is the
// 1e-6 jitter the callers put on the diagonal (src/cholesky-cache.cpp:22, src/draw-fstar.cpp:35).
__global__ void k_build_K(const double* __restrict__ x1, int n1, const double* __restrict__ x2, int n2,
                          double s0, double s1, double s2, double* __restrict__ out, int ld, double diag_add) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int j = blockIdx.y;
    if (i >= n1 || j >= n2) return;
    double a = x1[i], b = x2[j];
    double diff = __dsub_rn(a, b);
    double r = exp(__dmul_rn(__dmul_rn(-0.5, diff), diff));
    r = __dadd_rn(r, __dmul_rn(__dmul_rn(a, __dmul_rn(s1, s1)), b));
    r = __dadd_rn(r, __dmul_rn(s0, s0));
    double q = __dmul_rn(__dmul_rn(a, s2), b);
    r = __dadd_rn(r, __dmul_rn(q, q));
    if (i == j) r = __dadd_rn(r, diag_add);
    out[(size_t)j * ld + i] = r;
}

cudaError_t build_K(cudaStream_t st, const double* x1, int n1, const double* x2, int n2, const double* sds_host3,
                    double* out, int ld, double diag_add) {
    dim3 grid((n1 + 255) / 256, n2);
    k_build_K<<<grid, 256, 0, st>>>(x1, n1, x2, n2, sds_host3[0], sds_host3[1], sds_host3[2], out, ld, diag_add);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// K_time (src/covariance-function.cpp:21-44): Matern-5/2 (kernel 0) or RBF (kernel 1) over session indices.
__global__ void k_build_K_time(const double* __restrict__ x1, int n1, const double* __restrict__ x2, int n2, double os,
                               double ls, double t0, double t1, int kernel, double* __restrict__ out, int ld) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int j = blockIdx.y;
    if (i >= n1 || j >= n2) return;
    double a = x1[i], b = x2[j];
    double diff = fabs(__dsub_rn(a, b));
    double r = 0.0;
    const double s5 = sqrt(5.0);
    if (kernel == 0) {
        double t = __dadd_rn(__dadd_rn(1.0, __ddiv_rn(__dmul_rn(s5, diff), ls)),
                             __ddiv_rn(__ddiv_rn(__ddiv_rn(__dmul_rn(__dmul_rn(5.0, diff), diff), ls), ls), 3.0));
        r = __dmul_rn(__dmul_rn(os, os), t);
        r = __dmul_rn(r, exp(__ddiv_rn(__dmul_rn(-s5, diff), ls)));
        r = __dadd_rn(r, __dmul_rn(__dmul_rn(a, __dmul_rn(t1, t1)), b));
        r = __dadd_rn(r, __dmul_rn(t0, t0));
    } else if (kernel == 1) {
        r = __dmul_rn(__dmul_rn(os, os), exp(__ddiv_rn(__ddiv_rn(__dmul_rn(-diff, diff), ls), ls)));
        r = __dadd_rn(r, __dmul_rn(__dmul_rn(a, __dmul_rn(t1, t1)), b));
        r = __dadd_rn(r, __dmul_rn(t0, t0));
    }
    out[(size_t)j * ld + i] = r;
}

cudaError_t build_K_time(cudaStream_t st, const double* x1, int n1, const double* x2, int n2, double os, double ls,
                         double t0, double t1, int kernel, double* out, int ld) {
    dim3 grid((n1 + 255) / 256, n2);
    k_build_K_time<<<grid, 256, 0, st>>>(x1, n1, x2, n2, os, ls, t0, t1, kernel, out, ld);
    ++g_kernel_launches;
    return cudaGetLastError();
}

}  // namespace gpirt

namespace gpirt {

// ------------------------------------------------------------------------------------------------
// Cholesky of one NB x NB diagonal block, one CTA, block resident in shared memory.
__global__ void __launch_bounds__(256) k_potf2(double* __restrict__ A, int ld, int kb, int* info, int k0) {
    constexpr int NB = CHOL_NB;
    __shared__ double S[NB][NB + 1];
    const int tid = threadIdx.x;
    for (int e = tid; e < NB * NB; e += 256) {
        int r = e % NB, c = e / NB;
        S[r][c] = (r < kb && c < kb && r >= c) ? A[(size_t)c * ld + r] : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    for (int j = 0; j < kb; ++j) {
        if (tid == 0) {
            double d = S[j][j];
            if (!(d > 0.0)) {  // arma::chol would throw here; record the failing leading minor and keep going
                atomicCAS(info, 0, k0 + j + 1);
                d = 1.0;
            }
            S[j][j] = sqrt(d);
        }
        __syncthreads();
        const double d = S[j][j];
        for (int i = j + 1 + tid; i < kb; i += 256) S[i][j] = S[i][j] / d;
        __syncthreads();
        const int t = kb - j - 1;
        for (int e = tid; e < t * t; e += 256) {
            int r = j + 1 + e % t, c = j + 1 + e / t;
            if (r >= c) S[r][c] -= S[r][j] * S[c][j];
        }
        __syncthreads();
    }
    for (int e = tid; e < NB * NB; e += 256) {
        int r = e % NB, c = e / NB;
        if (r < kb && c < kb && r >= c) A[(size_t)c * ld + r] = S[r][c];
    }
}

// Panel solve X * L11^T = A21 (rows of the panel are independent): one thread per row keeps its 64
// unknowns in registers, L11 is broadcast from shared memory. Backward-stable substitution.
__global__ void __launch_bounds__(64) k_trsm_right_lt(double* __restrict__ A21, int ld, int rows,
                                                      const double* __restrict__ L11) {
    constexpr int NB = CHOL_NB;
    __shared__ double Ls[NB][NB + 1];
    for (int e = threadIdx.x; e < NB * NB; e += 64) {
        int r = e % NB, c = e / NB;
        Ls[r][c] = (r >= c) ? L11[(size_t)c * ld + r] : 0.0;
    }
    __syncthreads();
    const int r = blockIdx.x * 64 + threadIdx.x;
    if (r >= rows) return;
    double x[NB];
#pragma unroll
    for (int k = 0; k < NB; ++k) x[k] = A21[(size_t)k * ld + r];
#pragma unroll
    for (int k = 0; k < NB; ++k) {
        const double xk = x[k] / Ls[k][k];
        x[k] = xk;
#pragma unroll
        for (int i = k + 1; i < NB; ++i) x[i] -= xk * Ls[i][k];
    }
#pragma unroll
    for (int k = 0; k < NB; ++k) A21[(size_t)k * ld + r] = x[k];
}

// Diagonal-block solves for many right-hand sides: one thread per column of B.
template <bool TRANS>
__global__ void __launch_bounds__(64) k_trsm_left(const double* __restrict__ L11, int ldl, int kb,
                                                  double* __restrict__ B, int ldb, int nrhs) {
    constexpr int NB = CHOL_NB;
    __shared__ double Ls[NB][NB + 1];
    for (int e = threadIdx.x; e < NB * NB; e += 64) {
        int r = e % NB, c = e / NB;
        Ls[r][c] = (r < kb && c < kb) ? (r >= c ? L11[(size_t)c * ldl + r] : 0.0) : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    const int c = blockIdx.x * 64 + threadIdx.x;
    if (c >= nrhs) return;
    double* b = B + (size_t)c * ldb;
    double x[NB];
#pragma unroll
    for (int k = 0; k < NB; ++k) x[k] = (k < kb) ? b[k] : 0.0;
    if (!TRANS) {
#pragma unroll
        for (int k = 0; k < NB; ++k) {
            const double xk = x[k] / Ls[k][k];
            x[k] = xk;
#pragma unroll
            for (int i = k + 1; i < NB; ++i) x[i] -= xk * Ls[i][k];
        }
    } else {
#pragma unroll
        for (int k = NB - 1; k >= 0; --k) {
            const double xk = x[k] / Ls[k][k];
            x[k] = xk;
#pragma unroll
            for (int i = 0; i < k; ++i) x[i] -= xk * Ls[k][i];
        }
    }
#pragma unroll
    for (int k = 0; k < NB; ++k)
        if (k < kb) b[k] = x[k];
}

__global__ void k_zero_upper(double* __restrict__ A, int ld, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int j = blockIdx.y;
    if (i < n && j < n && i < j) A[(size_t)j * ld + i] = 0.0;
}

cudaError_t chol_lower_blocked(cudaStream_t st, double* A, int ld, int n, int* info_dev) {
    constexpr int NB = CHOL_NB;
    for (int k0 = 0; k0 < n; k0 += NB) {
        const int kb = (n - k0 < NB) ? n - k0 : NB;
        double* Akk = A + (size_t)k0 * ld + k0;
        k_potf2<<<1, 256, 0, st>>>(Akk, ld, kb, info_dev, k0);
        ++g_kernel_launches;
        const int rows = n - k0 - kb;
        if (rows > 0) {
            double* A21 = Akk + kb;
            k_trsm_right_lt<<<(rows + 63) / 64, 64, 0, st>>>(A21, ld, rows, Akk);
            ++g_kernel_launches;
            GemmArgs g{};
            g.M = rows; g.N = rows; g.K = kb;
            g.A = A21; g.lda = ld; g.B = A21; g.ldb = ld;
            g.C = A + (size_t)(k0 + kb) * ld + (k0 + kb); g.ldc = ld;
            g.alpha = -1.0; g.beta = 1.0; g.lower_only = 1;
            cudaError_t e = gemm_f64(st, false, true, g);
            if (e != cudaSuccess) return e;
        }
    }
    dim3 grid((n + 255) / 256, n);
    k_zero_upper<<<grid, 256, 0, st>>>(A, ld, n);
    ++g_kernel_launches;
    return cudaGetLastError();
}

cudaError_t trsm_left_lower(cudaStream_t st, const double* L, int ldl, int n, double* B, int ldb, int nrhs) {
    constexpr int NB = CHOL_NB;
    for (int k0 = 0; k0 < n; k0 += NB) {
        const int kb = (n - k0 < NB) ? n - k0 : NB;
        k_trsm_left<false><<<(nrhs + 63) / 64, 64, 0, st>>>(L + (size_t)k0 * ldl + k0, ldl, kb, B + k0, ldb, nrhs);
        ++g_kernel_launches;
        const int rows = n - k0 - kb;
        if (rows > 0) {
            GemmArgs g{};
            g.M = rows; g.N = nrhs; g.K = kb;
            g.A = L + (size_t)k0 * ldl + k0 + kb; g.lda = ldl;
            g.B = B + k0; g.ldb = ldb;
            g.C = B + k0 + kb; g.ldc = ldb;
            g.alpha = -1.0; g.beta = 1.0;
            cudaError_t e = gemm_f64(st, false, false, g);
            if (e != cudaSuccess) return e;
        }
    }
    return cudaGetLastError();
}

cudaError_t trsm_left_lower_t(cudaStream_t st, const double* L, int ldl, int n, double* B, int ldb, int nrhs) {
    constexpr int NB = CHOL_NB;
    const int last = ((n - 1) / NB) * NB;
    for (int k0 = last; k0 >= 0; k0 -= NB) {
        const int kb = (n - k0 < NB) ? n - k0 : NB;
        k_trsm_left<true><<<(nrhs + 63) / 64, 64, 0, st>>>(L + (size_t)k0 * ldl + k0, ldl, kb, B + k0, ldb, nrhs);
        ++g_kernel_launches;
        if (k0 > 0) {
            // B[0:k0, :] -= L[k0:k0+kb, 0:k0]^T * X_k
            GemmArgs g{};
            g.M = k0; g.N = nrhs; g.K = kb;
            g.A = L + k0; g.lda = ldl;
            g.B = B + k0; g.ldb = ldb;
            g.C = B; g.ldc = ldb;
            g.alpha = -1.0; g.beta = 1.0;
            cudaError_t e = gemm_f64(st, true, false, g);
            if (e != cudaSuccess) return e;
        }
    }
    return cudaGetLastError();
}

}  // namespace gpirt
