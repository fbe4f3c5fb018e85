// Dense fp64 building blocks of the sweep: covariance construction (src/covariance-function.cpp:3-44),
// blocked lower Cholesky (arma::chol(S,"lower"), src/cholesky-cache.cpp:23) and the triangular solves
// behind double_solve / solve(trimatl(L), kstar) (src/cholesky-cache.cpp:4-9, src/draw-fstar.cpp:30).
#include "linalg.cuh"

#include <atomic>
#include <cstdio>
#include <cstdlib>

namespace gpirt {

long long g_kernel_launches = 0;

// ------------------------------------------------------------------------------------------------
// GEMM dispatch
template <int BM, int BN, int BK, int WM, int WN, int STAGES, bool TA, bool TB>
static cudaError_t launch_gemm(cudaStream_t st, const GemmArgs& g) {
    using Cfg = GemmCfg<BM, BN, BK, WM, WN, STAGES, TA, TB>;
    auto kern = gemm_f64_kernel<BM, BN, BK, WM, WN, STAGES, TA, TB>;
    static std::atomic<unsigned long long> configured{0};  // per instantiation, one bit per device: the opt-in is a per-device attribute
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 64 || !(configured.load(std::memory_order_relaxed) >> dev & 1ULL)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        if (dev < 64) configured.fetch_or(1ULL << dev, std::memory_order_relaxed);
    }
    dim3 grid(((g.N + BN - 1) / BN) * ((g.M + BM - 1) / BM), 1, g.batch > 1 ? g.batch : 1);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(g);
    ++g_kernel_launches;
    return cudaGetLastError();
}

static int sm_count() {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    return sms;
}

// Tile shape: 64 x 64 x 16 CTA tiles, four warps of 32 x 32, three cp.async stages: 124 registers, 4 CTAs (16 warps) per SM. Measured
// on B200 over the three big products of a C3 sweep (tools/bench_gemm.cu, profiles/r01_gemm_tiles.txt) it beats every larger tile —
// 128 x 128 needs ~200 registers, so one CTA (8 warps) per SM, and stalls on its own latency — and ncu shows the DMMA pipe 90-91 %
// busy (math-pipe throttle is the top stall). GPIRT_GEMM_TILE=128 selects the old big-tile path for experiments.
template <bool TA, bool TB>
static cudaError_t gemm_pick(cudaStream_t st, const GemmArgs& g) {
    static const int force = getenv("GPIRT_GEMM_TILE") ? atoi(getenv("GPIRT_GEMM_TILE")) : 0;
    if (force != 128) return launch_gemm<64, 64, 16, 32, 32, 3, TA, TB>(st, g);
    const long tm = (g.M + 127) / 128, tn = (g.N + 127) / 128;
    long tiles128 = tm * tn * (g.batch > 1 ? g.batch : 1);
    if (g.lower_only) tiles128 = tiles128 / 2 + 1;
    const int slots = sm_count();  // one 128 x 128 CTA per SM
    if (!(tiles128 >= 2 * slots && g.N >= 128 && g.M >= 128)) return launch_gemm<64, 64, 16, 32, 32, 3, TA, TB>(st, g);
    // wave quantisation: the rows of a mostly empty last wave go to the small-tile kernel
    const bool uniform = !g.a_lower && !g.a_upper && !g.b_lower && !g.b_upper && !g.lower_only && g.batch <= 1;
    const long full = tiles128 / slots, rem = tiles128 % slots;
    if (uniform && rem > 0 && rem * 10 < slots * 6) {
        const long rows_main = (full * slots) / tn;
        if (rows_main >= 1 && rows_main < tm) {
            GemmArgs a = g, b = g;
            a.M = (int)rows_main * 128;
            b.M = g.M - a.M;
            b.m_off = g.m_off + a.M;
            b.A = TA ? g.A + (size_t)a.M * g.lda : g.A + a.M;
            b.C = g.C + a.M;
            cudaError_t e = launch_gemm<128, 128, 16, 64, 32, 3, TA, TB>(st, a);
            if (e != cudaSuccess) return e;
            return launch_gemm<64, 64, 16, 32, 32, 3, TA, TB>(st, b);
        }
    }
    return launch_gemm<128, 128, 16, 64, 32, 3, TA, TB>(st, g);
}

cudaError_t gemm_f64(cudaStream_t st, bool ta, bool tb, const GemmArgs& g0) {
    if (g0.M <= 0 || g0.N <= 0) return cudaSuccess;
    GemmArgs g = g0;
    if (g.band_n <= 0) {  // a band's slice of op(B), 64 columns x K doubles per column tile, should stay in the L2 next to op(A) and the output. 41 MB = two bands at C3 (narrower bands measured slower: profiles/r02_sweep_schedule_experiments.txt)
        static const int band_mb = getenv("GPIRT_GEMM_BAND_MB") ? atoi(getenv("GPIRT_GEMM_BAND_MB")) : 41;
        const long long per_tile = (long long)(g.K > 0 ? g.K : 1) * 64 * 8;
        const long long bw = ((long long)band_mb << 20) / per_tile;
        g.band_n = (int)(bw < 4 ? 4 : bw);  // band_mb = 0 in the environment: one band (the order before)
        if (band_mb <= 0) g.band_n = 1 << 30;
    }
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
// 64 x 64 tile kernels: 256 threads as a 16 x 16 grid, thread (tr, tc) keeps the CONTIGUOUS 4 x 4 block rows 4tr..4tr+3,
// columns 4tc..4tc+3 of the tile in registers. Eliminations advance four columns (or rows) per step: the thread that owns the
// diagonal 4 x 4 block factors / inverts it locally, the 16 owners of the block column solve their 4 x 4 blocks against it, and
// everybody applies a rank-4 update — 16 steps of two barriers instead of 64 steps of one, and the sqrt / reciprocal chains
// (MUFU.RSQ64H + Newton; no fp64 division, which costs ~1500 cycles in divergent code, tools/bench_tile.cu) sit in one thread.
// (History: a fully unrolled 64-step substitution was instruction-fetch bound at 70-150 us per launch; the one-column-per-barrier
// version that followed spent ~600 cycles per column, profiles/r01_launches_v3.csv.)
constexpr int TILE_THREADS = 256;
#ifdef GPIRT_PANEL_CLOCKS
__device__ unsigned long long g_potf2_clk[4];
#define TCLK_BEGIN long long tclk_ = clock64()
#define TCLK_END(i, cond) do { if ((cond) && blockIdx.x == 1) atomicAdd(&g_potf2_clk[i], (unsigned long long)(clock64() - tclk_)); } while (0)
#else
#define TCLK_BEGIN
#define TCLK_END(i, cond)
#endif

struct TileShared {       // state of one elimination step; kernels keep two and alternate (one barrier less per step)
    double D[4][4];        // factored diagonal 4 x 4 block (lower part valid)
    double Dinv[4];        // reciprocals of its diagonal
    double P[4][CHOL_NB];  // the current block column, k-major: a thread's 4 rows are 32 contiguous bytes (two conflict-free LDS.128)
};

// sqrt(d) and 1/sqrt(d) without a division; s is corrected once so that it is the correctly rounded root up to 1 ulp.
__device__ __forceinline__ void sqrt_rsqrt(double d, double& s, double& r) {
    r = rsqrt(d);
    s = d * r;
    s = fma(fma(-s, s, d), 0.5 * r, s);
}

// X * D^T = A for one 4 x 4 block held by rows: x[q] <- (a[q] - sum_{k<q} x[k] D[q][k]) * Dinv[q]
__device__ __forceinline__ void row_solve4(double (&x)[4], const double (&D)[4][4], const double (&Dinv)[4]) {
    x[0] = x[0] * Dinv[0];
    x[1] = (x[1] - x[0] * D[1][0]) * Dinv[1];
    x[2] = (x[2] - x[0] * D[2][0] - x[1] * D[2][1]) * Dinv[2];
    x[3] = (x[3] - x[0] * D[3][0] - x[1] * D[3][1] - x[2] * D[3][2]) * Dinv[3];
}

// One elimination step s of X * L^T = A restricted to block column s (shared by the Cholesky and its panel solve): owners
// (tc == s, row block >= first_rb) solve against the 4 x 4 block D and publish their rows in P; after the barrier the columns to
// the right take the rank-4 update a[p][q] -= sum_k P[row][k] * Lcol(4tc+q, 4s+k), Lcol read from P itself (Cholesky: the
// panel is the factor; only tr >= tc matters) or from Ls (panel solve). One barrier; P is double-buffered by the caller.
template <bool FROM_LS>
__device__ __forceinline__ void step_block_column(double (&a)[4][4], int s, int tr, int tc, const double (&D)[4][4], const double (&Dinv)[4],
                                                  double (*P)[CHOL_NB], double (*Ls)[CHOL_NB + 1], int first_rb) {
    if (tc == s && tr >= first_rb) {
        TCLK_BEGIN;
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            row_solve4(a[p], D, Dinv);
#pragma unroll
            for (int q = 0; q < 4; ++q) P[q][4 * tr + p] = a[p][q];
        }
        TCLK_END(1, !FROM_LS && tr == 15);
    }
    __syncthreads();
    if (tc > s && (FROM_LS || tr >= tc)) {
        TCLK_BEGIN;
        double xr[4][4], lc[4][4];
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                xr[p][k] = P[k][4 * tr + p];
                lc[p][k] = FROM_LS ? Ls[4 * tc + p][4 * s + k] : P[k][4 * tc + p];
            }
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int k = 0; k < 4; ++k) a[p][q] -= xr[p][k] * lc[q][k];
        TCLK_END(2, !FROM_LS && tr == 15 && tc == 15);
    }
}

// In-register Cholesky of the tile held as a[4][4] per thread; returns with the factor in a (lower part valid).
__device__ __forceinline__ void tile_potf2(double (&a)[4][4], int tr, int tc, TileShared* sh2, int* info, int k0, bool report) {
    for (int s = 0; s < 16; ++s) {
        TileShared& sh = sh2[s & 1];
        if (tr == s && tc == s) {  // the diagonal 4 x 4 block: unblocked Cholesky in registers
            TCLK_BEGIN;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                double d = a[j][j];
                if (!(d > 0.0)) {  // arma::chol would throw here; record the failing leading minor and keep going
                    if (report) atomicCAS(info, 0, k0 + 4 * s + j + 1);
                    d = 1.0;
                }
                double sq, rs;
                sqrt_rsqrt(d, sq, rs);
                a[j][j] = sq;
                sh.Dinv[j] = rs;
#pragma unroll
                for (int i = j + 1; i < 4; ++i) a[i][j] *= rs;
#pragma unroll
                for (int jj = j + 1; jj < 4; ++jj)
#pragma unroll
                    for (int i = jj; i < 4; ++i) a[i][jj] -= a[i][j] * a[jj][j];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (j > i) a[i][j] = 0.0;
                    sh.D[i][j] = a[i][j];
                    sh.P[j][4 * s + i] = a[i][j];  // the factor's own rows of the block column (upper part zero)
                }
            TCLK_END(0, true);
        }
        __syncthreads();
        double D[4][4], Dinv[4];
        if (tc == s && tr > s) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Dinv[i] = sh.Dinv[i];
#pragma unroll
                for (int j = 0; j < 4; ++j) D[i][j] = sh.D[i][j];
            }
        }
        step_block_column<false>(a, s, tr, tc, D, Dinv, sh.P, nullptr, s + 1);
    }
}

// One panel of the blocked factorisation in ONE launch: every CTA factors the diagonal block (redundantly, bit-identical), CTA 0
// writes it back, CTA b >= 1 solves X * L11^T = A21 for its 64 rows of the panel by substitution (backward stable).
// In-place hazard: CTAs b >= 1 read the ORIGINAL diagonal block that CTA 0 overwrites with its factor. A CTA may start late (the
// panel runs next to other streams' kernels), so CTA 0 holds its write-back until every other CTA has announced, through *readers,
// that its copy of the block is in registers.
#ifdef GPIRT_PANEL_CLOCKS
__device__ long long g_panel_clk[8];
#define PCLK(i) do { __syncthreads(); if (blockIdx.x == 1 && threadIdx.x == 0) g_panel_clk[i] = clock64(); } while (0)
#else
#define PCLK(i)
#endif
__global__ void __launch_bounds__(TILE_THREADS) k_chol_panel(double* __restrict__ A, int ld, int kb, int rows_below, int* info, int k0,
                                                             int* readers) {
    constexpr int NB = CHOL_NB;
    __shared__ double Ls[NB][NB + 1];
    __shared__ TileShared sh2[2];
    __shared__ double inv_diag[NB];
    const int tr = threadIdx.x & 15, tc = threadIdx.x >> 4;
    double a[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = 4 * tr + p, c = 4 * tc + q;
            a[p][q] = (r < kb && c < kb && c <= r) ? A[(size_t)c * ld + r] : (c == r ? 1.0 : 0.0);
        }
    PCLK(0);
    if (blockIdx.x != 0) {
        __threadfence();  // this thread's loads of the block are performed
        __syncthreads();
        if (threadIdx.x == 0) atomicAdd(readers, 1);
    }
    PCLK(1);
    tile_potf2(a, tr, tc, sh2, info, k0, blockIdx.x == 0);
    PCLK(2);
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) {
            while (atomicAdd(readers, 0) < (int)gridDim.x - 1) __nanosleep(64);
            __threadfence();
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const int r = 4 * tr + p, c = 4 * tc + q;
                if (r < kb && c < kb) A[(size_t)c * ld + r] = (c <= r) ? a[p][q] : 0.0;  // the block's upper part is zeroed here: consumers of leading rows
                                                                                          // (CholMilestones) read the diagonal block before k_zero_upper runs
            }
        return;
    }
    __syncthreads();  // the last update of tile_potf2 still reads sh2
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = 4 * tr + p, c = 4 * tc + q;
            Ls[r][c] = (c <= r) ? a[p][q] : 0.0;
        }
    if (tr == tc) {
#pragma unroll
        for (int p = 0; p < 4; ++p) inv_diag[4 * tr + p] = rsqrt(a[p][p] * a[p][p]);  // 1 / L(jj): no fp64 division
    }
    // X * L11^T = A21 for rows [row0, row0 + 64) of the panel; kb == NB whenever rows_below > 0
    const int row0 = (blockIdx.x - 1) * 64;
    double* __restrict__ A21 = A + kb;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = row0 + 4 * tr + p, c = 4 * tc + q;
            a[p][q] = (r < rows_below) ? A21[(size_t)c * ld + r] : 0.0;
        }
    __syncthreads();  // Ls, inv_diag visible
    PCLK(3);
    for (int s = 0; s < 16; ++s) {
        double D[4][4], Dinv[4];
        if (tc == s) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Dinv[i] = inv_diag[4 * s + i];
#pragma unroll
                for (int j = 0; j < 4; ++j) D[i][j] = Ls[4 * s + i][4 * s + j];
            }
        }
        step_block_column<true>(a, s, tr, tc, D, Dinv, sh2[s & 1].P, Ls, 0);
    }
    PCLK(4);
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = row0 + 4 * tr + p, c = 4 * tc + q;
            if (r < rows_below) A21[(size_t)c * ld + r] = a[p][q];
        }
    PCLK(5);
}

// Substitution with a 64 x 64 triangular block held in shared memory as M[i][j] = L(j, i) (TRANS false) or L(i, j) (TRANS true), so
// that the coefficients a thread needs for its 4 rows are contiguous: on entry a[p][q] = B(row 4tr + p, column 4tc + q), on exit
// the solution of L X = B or L^T X = B. Four rows per step: the 16 owners of the row block solve their 4 x 4 blocks against the
// diagonal block, publish them in XR, and the remaining row blocks take a rank-4 update. One barrier per step (XR double-buffered).
template <bool TRANS>
__device__ __forceinline__ void tile_solve(double (&a)[4][4], double (*M)[CHOL_NB], double (*XR2)[4][CHOL_NB], const double* inv_diag, int tr, int tc) {
    for (int t = 0; t < 16; ++t) {
        const int s = TRANS ? 15 - t : t;
        double (*XR)[CHOL_NB] = XR2[t & 1];
        if (tr == s) {
            double D[4][4], Dinv[4];  // D[i][j] = L(4s+i, 4s+j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Dinv[i] = inv_diag[4 * s + i];
#pragma unroll
                for (int j = 0; j < 4; ++j) D[i][j] = TRANS ? M[4 * s + i][4 * s + j] : M[4 * s + j][4 * s + i];
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (!TRANS) {
                    a[0][q] = a[0][q] * Dinv[0];
                    a[1][q] = (a[1][q] - D[1][0] * a[0][q]) * Dinv[1];
                    a[2][q] = (a[2][q] - D[2][0] * a[0][q] - D[2][1] * a[1][q]) * Dinv[2];
                    a[3][q] = (a[3][q] - D[3][0] * a[0][q] - D[3][1] * a[1][q] - D[3][2] * a[2][q]) * Dinv[3];
                } else {  // D^T x = b: back substitution, D^T(i, j) = D[j][i]
                    a[3][q] = a[3][q] * Dinv[3];
                    a[2][q] = (a[2][q] - D[3][2] * a[3][q]) * Dinv[2];
                    a[1][q] = (a[1][q] - D[3][1] * a[3][q] - D[2][1] * a[2][q]) * Dinv[1];
                    a[0][q] = (a[0][q] - D[3][0] * a[3][q] - D[2][0] * a[2][q] - D[1][0] * a[1][q]) * Dinv[0];
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) XR[k][4 * tc + q] = a[k][q];
            }
        }
        __syncthreads();
        if (TRANS ? (tr < s) : (tr > s)) {
            double l[4][4], x[4][4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    l[p][k] = M[4 * s + k][4 * tr + p];  // coefficient of x_(4s+k) in the equation of row 4tr+p
                    x[k][p] = XR[k][4 * tc + p];
                }
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int k = 0; k < 4; ++k) a[p][q] -= l[p][k] * x[k][q];
        }
    }
}

// Diagonal-block solves for many right-hand sides: CTA b solves L11 X = B (or L11^T X = B) for 64 columns of B.
template <bool TRANS>
__global__ void __launch_bounds__(TILE_THREADS) k_trsm_left(const double* __restrict__ L11, int ldl, int kb,
                                                            double* __restrict__ B, int ldb, int nrhs) {
    constexpr int NB = CHOL_NB;
    __shared__ __align__(16) double M[NB][NB];
    __shared__ __align__(16) double XR2[2][4][NB];
    __shared__ double inv_diag[NB];
    const int tr = threadIdx.x & 15, tc = threadIdx.x >> 4;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = 4 * tr + p, c = 4 * tc + q;
            const double v = (r < kb && c < kb) ? (c <= r ? L11[(size_t)c * ldl + r] : 0.0) : (r == c ? 1.0 : 0.0);
            if (TRANS) M[r][c] = v; else M[c][r] = v;
            if (r == c) inv_diag[r] = rsqrt(v * v);  // 1 / L(r, r) without an fp64 division
        }
    const int c0 = blockIdx.x * 64;
    double a[4][4];  // a[p][q] = X(row 4tr + p, column c0 + 4tc + q)
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = 4 * tr + p, c = c0 + 4 * tc + q;
            a[p][q] = (r < kb && c < nrhs) ? B[(size_t)c * ldb + r] : 0.0;
        }
    __syncthreads();
    tile_solve<TRANS>(a, M, XR2, inv_diag, tr, tc);
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = 4 * tr + p, c = c0 + 4 * tc + q;
            if (r < kb && c < nrhs) B[(size_t)c * ldb + r] = a[p][q];
        }
}

// Inverses of the 64 x 64 diagonal blocks of a lower-triangular L, one CTA per block: L_kk X = I by substitution.
// Writes the whole block of Linv (zeros above the diagonal).
__global__ void __launch_bounds__(TILE_THREADS) k_trtri_diag(const double* __restrict__ L, int ld, int n, double* __restrict__ Linv, int ldi) {
    constexpr int NB = CHOL_NB;
    __shared__ __align__(16) double M[NB][NB];
    __shared__ __align__(16) double XR2[2][4][NB];
    __shared__ double inv_diag[NB];
    const int tr = threadIdx.x & 15, tc = threadIdx.x >> 4;
    const int k0 = blockIdx.x * NB;
    const int kb = min(NB, n - k0);
    const double* L11 = L + (size_t)k0 * ld + k0;
    double a[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = 4 * tr + p, c = 4 * tc + q;
            const double v = (r < kb && c < kb) ? (c <= r ? L11[(size_t)c * ld + r] : 0.0) : (r == c ? 1.0 : 0.0);
            M[c][r] = v;
            if (r == c) inv_diag[r] = rsqrt(v * v);
            a[p][q] = (r == c) ? 1.0 : 0.0;
        }
    __syncthreads();
    tile_solve<false>(a, M, XR2, inv_diag, tr, tc);
    double* out = Linv + (size_t)k0 * ldi + k0;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = 4 * tr + p, c = 4 * tc + q;
            if (r < kb && c < kb) out[(size_t)c * ldi + r] = (c <= r) ? a[p][q] : 0.0;
        }
}

__global__ void k_zero_upper(double* __restrict__ A, int ld, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int j = blockIdx.y;
    if (i < n && j < n && i < j) A[(size_t)j * ld + i] = 0.0;
}

bool chol_is_persistent(int n, const CholPlan* plan) {
    static const int persistent_env = getenv("GPIRT_CHOL_PERSISTENT") ? atoi(getenv("GPIRT_CHOL_PERSISTENT")) : 1;
    return plan && plan->state && persistent_env && n < 4096;
}

cudaError_t chol_lower_blocked(cudaStream_t st, double* A, int ld, int n, int* info_dev, int* readers_ws, const CholMilestones* marks,
                               const CholAux* aux, const CholPlan* plan, const CholDist* dist) {
    if (chol_is_persistent(n, plan)) {
        cudaError_t e = chol_lower_persistent(st, A, ld, n, info_dev, plan->state, plan->tasks, plan->n_tasks, plan->max_ctas);
        if (e != cudaSuccess) return e;
        for (int c = 0; marks && c < marks->count; ++c)
            if ((e = cudaEventRecord(marks->ev[c], st)) != cudaSuccess) return e;
        return cudaSuccess;
    }
    int next_mark = 0;
    static const int lookahead_env = getenv("GPIRT_CHOL_LOOKAHEAD") ? atoi(getenv("GPIRT_CHOL_LOOKAHEAD")) : 1;
    bool rest_pending = false;  // a part of a trailing update is in flight on the helper stream
    constexpr int NB = CHOL_NB;
    cudaError_t e0 = cudaMemsetAsync(readers_ws, 0, chol_ws_ints(n) * sizeof(int), st);
    if (e0 != cudaSuccess) return e0;
    // Two levels of blocking for large matrices: 64-column panels (the serial part) inside OB-column outer blocks. After a panel only the
    // rest of its outer block is updated (a narrow K = 64 product); the trailing matrix takes one K = OB update per outer block, where the
    // DMMA GEMM runs near its peak. At n = 2000 the K = 64 updates are latency-bound launches either way (no gain, profiles/
    // r01_gemm_tiles.txt: OB = NB there); at n = 10 000 (config C4) the flops dominate and the K = 64 updates reach only ~16 TFLOP/s.
    static const int ob_env = getenv("GPIRT_CHOL_OB") ? atoi(getenv("GPIRT_CHOL_OB")) : 0;
    const bool distributed = dist && dist->world > 1 && dist->push && dist->wait && n >= 4096;
    const int OB = distributed ? 256 : (ob_env > 0 ? ob_env : (n >= 4096 ? 256 : NB));
    for (int o0 = 0; o0 < n; o0 += OB) {
        const int ob = (n - o0 < OB) ? n - o0 : OB;
        if (distributed && o0 > 0) {  // columns [o0, o0 + ob), rows o0 .. n-1, as updated by their owner
            const int b = o0 / OB, owner = chol_dist_owner(b, dist->world);
            if (owner != dist->rank) {
                cudaError_t ew = dist->wait(dist->ctx, st, b, owner);
                if (ew != cudaSuccess) return ew;
            }
        }
        for (int k0 = o0; k0 < o0 + ob; k0 += NB) {
            const int kb = (n - k0 < NB) ? n - k0 : NB;
            double* Akk = A + (size_t)k0 * ld + k0;
            const int rows = n - k0 - kb;
            k_chol_panel<<<1 + (rows + 63) / 64, TILE_THREADS, 0, st>>>(Akk, ld, kb, rows, info_dev, k0, readers_ws + k0 / NB);
            ++g_kernel_launches;
            while (marks && next_mark < marks->count && marks->rows[next_mark] <= k0 + kb) {  // rows [0, k0 + kb) of L are final now
                cudaError_t em = cudaEventRecord(marks->ev[next_mark], st);
                if (em != cudaSuccess) return em;
                ++next_mark;
            }
            const int wcols = (OB == NB) ? rows : o0 + ob - (k0 + kb);  // columns updated now: everything (one level) or the rest of the outer block
            if (rows > 0 && wcols > 0) {
                double* A21 = Akk + kb;
                GemmArgs g{};
                g.M = rows; g.N = wcols; g.K = kb;
                g.A = A21; g.lda = ld; g.B = A21; g.ldb = ld;
                g.C = A + (size_t)(k0 + kb) * ld + (k0 + kb); g.ldc = ld;
                g.alpha = -1.0; g.beta = 1.0; g.lower_only = 1;
                cudaError_t e = gemm_f64(st, false, true, g);
                if (e != cudaSuccess) return e;
            }
        }
        const int rows_t = n - o0 - ob;
        if (distributed && rows_t > 0) {
            double* P = A + (size_t)o0 * ld + (o0 + ob);  // rows below the outer block, its ob columns
            const bool ahead = aux && aux->helper && lookahead_env;
            cudaError_t e;
            auto update_block = [&](cudaStream_t s, int c0) {  // column block (o0 + ob + c0) / OB of the trailing matrix, from its diagonal down
                const int w = (rows_t - c0 < OB) ? rows_t - c0 : OB;
                GemmArgs g{};
                g.M = rows_t - c0; g.N = w; g.K = ob;
                g.A = P + c0; g.lda = ld; g.B = P + c0; g.ldb = ld;
                g.C = A + (size_t)(o0 + ob + c0) * ld + (o0 + ob + c0); g.ldc = ld;
                g.alpha = -1.0; g.beta = 1.0; g.lower_only = 1;
                return gemm_f64(s, false, true, g);
            };
            if (ahead && (e = cudaEventRecord(aux->ev_panel, st)) != cudaSuccess) return e;
            const int b_next = (o0 + ob) / OB;
            if (chol_dist_owner(b_next, dist->world) == dist->rank) {  // the next outer block is this rank's: updated first, on the panels' stream, and published
                if (rest_pending) {  // its earlier updates ran on the helper stream
                    if ((e = cudaStreamWaitEvent(st, aux->ev_rest, 0)) != cudaSuccess) return e;
                    rest_pending = false;
                }
                if ((e = update_block(st, 0)) != cudaSuccess) return e;
                const int w = rows_t < OB ? rows_t : OB;
                if ((e = dist->push(dist->ctx, st, b_next, A + (size_t)(o0 + ob) * ld + (o0 + ob), ld, rows_t, w)) != cudaSuccess) return e;
            }
            // this rank's other column blocks: under the next outer block's panels (helper stream), which only need the block just published
            cudaStream_t s2 = ahead ? aux->helper : st;
            bool any = false;
            for (int c0 = OB; c0 < rows_t; c0 += OB) {
                if (chol_dist_owner((o0 + ob + c0) / OB, dist->world) != dist->rank) continue;
                if (!any && ahead && (e = cudaStreamWaitEvent(s2, aux->ev_panel, 0)) != cudaSuccess) return e;
                any = true;
                if ((e = update_block(s2, c0)) != cudaSuccess) return e;
            }
            if (any && ahead) {
                if ((e = cudaEventRecord(aux->ev_rest, s2)) != cudaSuccess) return e;
                rest_pending = true;
            }
            continue;
        }
        if (OB != NB && rows_t > 0) {
            double* P = A + (size_t)o0 * ld + (o0 + ob);  // rows below the outer block, its ob columns
            const bool split = aux && aux->helper && lookahead_env && rows_t > OB;
            cudaError_t e;
            if (split) {  // the outer block's panels are done: the helper stream may start on the columns beyond the next outer block
                if ((e = cudaEventRecord(aux->ev_panel, st)) != cudaSuccess) return e;
            }
            if (rest_pending) {  // the previous block's share on the helper stream wrote the entries this update adds to
                if ((e = cudaStreamWaitEvent(st, aux->ev_rest, 0)) != cudaSuccess) return e;
                rest_pending = false;
            }
            GemmArgs g{};
            g.M = rows_t; g.N = split ? OB : rows_t; g.K = ob;
            g.A = P; g.lda = ld; g.B = P; g.ldb = ld;
            g.C = A + (size_t)(o0 + ob) * ld + (o0 + ob); g.ldc = ld;
            g.alpha = -1.0; g.beta = 1.0; g.lower_only = 1;
            if ((e = gemm_f64(st, false, true, g)) != cudaSuccess) return e;
            if (split) {
                if ((e = cudaStreamWaitEvent(aux->helper, aux->ev_panel, 0)) != cudaSuccess) return e;
                GemmArgs r{};
                r.M = rows_t - OB; r.N = rows_t - OB; r.K = ob;
                r.A = P + OB; r.lda = ld; r.B = P + OB; r.ldb = ld;
                r.C = A + (size_t)(o0 + ob + OB) * ld + (o0 + ob + OB); r.ldc = ld;
                r.alpha = -1.0; r.beta = 1.0; r.lower_only = 1;
                if ((e = gemm_f64(aux->helper, false, true, r)) != cudaSuccess) return e;
                if ((e = cudaEventRecord(aux->ev_rest, aux->helper)) != cudaSuccess) return e;
                rest_pending = true;
            }
        }
    }
    if (rest_pending) {
        cudaError_t e = cudaStreamWaitEvent(st, aux->ev_rest, 0);
        if (e != cudaSuccess) return e;
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
        k_trsm_left<false><<<(nrhs + 63) / 64, TILE_THREADS, 0, st>>>(L + (size_t)k0 * ldl + k0, ldl, kb, B + k0, ldb, nrhs);
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
        k_trsm_left<true><<<(nrhs + 63) / 64, TILE_THREADS, 0, st>>>(L + (size_t)k0 * ldl + k0, ldl, kb, B + k0, ldb, nrhs);
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

size_t trtri_scratch_elems(int n) {
    size_t need = 1;
    for (int b = CHOL_NB; b < n; b *= 2) {
        const size_t pairs = ((size_t)n + 2 * b - 1) / (2 * b);
        need = need > pairs * b * b ? need : pairs * b * b;
    }
    return need;
}

// Linv = L^-1 for lower-triangular L. Diagonal 64-blocks by substitution (one launch, all blocks in parallel), then log2(n/64)
// merge levels: inv([A 0; C B]) = [A^-1 0; -B^-1 C A^-1  B^-1], each level two batched GEMMs over all pairs. No sequential chain
// over the n/64 block rows, unlike a blocked TRSM: 1 + 2*levels launches in total. Linv's strict upper triangle must be zero on
// entry (it is never written).
cudaError_t trtri_lower(cudaStream_t st, const double* L, int ld, int n, double* Linv, int ldi, double* scratch, const TrtriDist* dist) {
    constexpr int NB = CHOL_NB;
    k_trtri_diag<<<(n + NB - 1) / NB, TILE_THREADS, 0, st>>>(L, ld, n, Linv, ldi);
    ++g_kernel_launches;
    for (int b = NB; b < n; b *= 2) {
        // pair z covers rows/cols [z*2b, z*2b + 2b); it exists when its second half is not empty
        int pairs = 0, last_b2 = 0;
        for (int r0 = 0; r0 + b < n; r0 += 2 * b) { ++pairs; last_b2 = (n - r0 - b < b) ? n - r0 - b : b; }
        if (pairs == 0) break;
        // this rank's columns [c_lo, c_hi) of every pair's b x b block. Work per column at relative position x: 2(1 - x) in the first product (A^-1 is
        // lower triangular) + 1 in the second: the cumulative 3x - x^2 is cut into `world` equal parts.
        const bool split = dist && dist->push && dist->sync && trtri_dist_level(b, dist->world);
        int c_lo = 0, c_hi = b;
        if (split) {
            c_lo = trtri_dist_cut(b, dist->world, dist->rank);
            c_hi = trtri_dist_cut(b, dist->world, dist->rank + 1);
        }
        cudaError_t e;
        if (c_hi > c_lo) {
            GemmArgs g{};  // T_z = C_z * A_z^-1   (C_z = L[r0+b.., r0..], A_z^-1 lower triangular)
            g.M = b; g.N = c_hi - c_lo; g.K = b; g.n_off = c_lo;
            g.A = L + b; g.lda = ld; g.strideA = (long long)2 * b * ((long long)ld + 1);
            g.B = Linv + (size_t)c_lo * ldi; g.ldb = ldi; g.strideB = (long long)2 * b * ((long long)ldi + 1); g.b_lower = 1;
            g.C = scratch + (size_t)c_lo * b; g.ldc = b; g.strideC = (long long)b * b;
            g.alpha = 1.0; g.beta = 0.0; g.batch = pairs > 1 ? pairs : 0; g.M_last = last_b2;
            if (pairs == 1) g.M = last_b2;
            if ((e = gemm_f64(st, false, false, g)) != cudaSuccess) return e;
            GemmArgs h{};  // Linv[r0+b.., r0..] = -B_z^-1 * T_z   (B_z^-1 lower triangular)
            h.M = b; h.N = c_hi - c_lo; h.K = b;
            h.A = Linv + (size_t)b * ldi + b; h.lda = ldi; h.strideA = (long long)2 * b * ((long long)ldi + 1); h.a_lower = 1;
            h.B = scratch + (size_t)c_lo * b; h.ldb = b; h.strideB = (long long)b * b;
            h.C = Linv + b + (size_t)c_lo * ldi; h.ldc = ldi; h.strideC = (long long)2 * b * ((long long)ldi + 1);
            h.alpha = -1.0; h.beta = 0.0; h.batch = pairs > 1 ? pairs : 0; h.M_last = last_b2; h.K_last = last_b2;
            if (pairs == 1) { h.M = last_b2; h.K = last_b2; }
            if ((e = gemm_f64(st, false, false, h)) != cudaSuccess) return e;
            for (int z = 0; split && z < pairs; ++z) {
                double* blk = Linv + b + (size_t)z * 2 * b * ((size_t)ldi + 1) + (size_t)c_lo * ldi;
                if ((e = dist->push(dist->ctx, st, blk, ldi, z == pairs - 1 ? last_b2 : b, c_hi - c_lo)) != cudaSuccess) return e;
            }
        }
        if (split && (e = dist->sync(dist->ctx, st)) != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

}  // namespace gpirt
