// FP64 GEMM on the DMMA.8x8x4 pipe (mma.sync.m8n8k4.f64), cp.async multi-stage shared-memory pipeline.
// C(MxN) = alpha * op(A) * op(B) + beta * C, everything column-major.
//
// Measured on this pool's B200 (profiles/r01_microbench_fp64.txt): DMMA.8x8x4 37.1 TFLOP/s, scalar DFMA
// 34.0, cuBLAS DGEMM 8192^3 35.5 — so the FP64 tensor pipe is the roofline for every dense contraction of
// the sweep (L*Z prior draws, Cholesky trailing updates, the f* mean and noise products). tcgen05 has no
// fp64 kind, so mma.sync DMMA is the B200 fp64 tensor path.
//
// Shared-memory layouts are picked per operand so the global->shared copy is always a straight 16-byte
// cp.async along the operand's contiguous dimension, and the DMMA fragment reads are bank-conflict free
// (row pitch == 4 mod 16 doubles):
//   A not transposed (M contiguous): As[k][m], pitch BM+4      A transposed (K contiguous): As[m][k], pitch BK+4
//   B transposed     (N contiguous): Bs[k][n], pitch BN+4      B not transposed (K contig.): Bs[n][k], pitch BK+4
// Requirements: lda/ldb even and operand base pointers 16-byte aligned (all buffers of the sampler are
// allocated with leading dimensions padded to 16 doubles; sub-blocks start at multiples of 64 rows).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gpirt {

// number of kernels this library has launched (bench.py reports it as gpu_launches)
extern long long g_kernel_launches;

struct GemmArgs {
    int M, N, K;
    const double* A; int lda;
    const double* B; int ldb;
    double* C; int ldc;
    double alpha, beta;
    int a_lower;      // op(A)[m][k] == 0 for k > m  (k loop of a row block stops at its last row)
    int b_upper;      // op(B)[k][n] == 0 for k > n  (k loop of a column block stops at its last column)
    int lower_only;   // skip output tiles strictly above the diagonal (symmetric updates)
    int a_upper;      // op(A)[m][k] == 0 for k < m  (k loop of a row block starts at its first row)
    int b_lower;      // op(B)[k][n] == 0 for k < n  (k loop of a column block starts at its first column)
    // batching over blockIdx.z: operand z sits at base + z * stride (elements); the last batch may be ragged
    int batch;        // 0 or 1: a single product
    long long strideA, strideB, strideC;
    int M_last, N_last, K_last;  // extents of batch (batch-1) when > 0, else M / N / K
    int band_n;       // column tiles per band of the tile order (0: chosen by gemm_f64 from K; only without b_upper / b_lower)
    int n_off;        // column index of C(0,0) inside the full product (column slices with b_upper / b_lower)
    int m_off;        // row index of C(0,0) inside the full product (a launch over a row slice keeps the triangular k ranges right)
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int BM, int BN, int BK, int WM, int WN, int STAGES, bool TA, bool TB>
struct GemmCfg {
    static constexpr int WARPS_M = BM / WM, WARPS_N = BN / WN;
    static constexpr int THREADS = WARPS_M * WARPS_N * 32;
    static constexpr int A_PITCH = TA ? BK + 4 : BM + 4;
    static constexpr int A_ROWS = TA ? BM : BK;
    static constexpr int B_PITCH = TB ? BN + 4 : BK + 4;
    static constexpr int B_ROWS = TB ? BK : BN;
    static constexpr int A_ELEMS = A_PITCH * A_ROWS, B_ELEMS = B_PITCH * B_ROWS;
    static constexpr int SMEM_BYTES = STAGES * (A_ELEMS + B_ELEMS) * 8;
};

// Copy one operand tile: `rows` rows of `cols` contiguous doubles (cols even), zero-filled outside
// [0,row_ext) x [0,col_ext). src points at element (0,0) of the tile; row stride ld (doubles).
template <int ROWS, int COLS, int PITCH, int THREADS>
__device__ __forceinline__ void load_tile(double* sm, const double* src, int ld, int row_ext, int col_ext, int tid) {
    constexpr int CHUNKS_PER_ROW = COLS / 2;
    constexpr int CHUNKS = ROWS * CHUNKS_PER_ROW;
#pragma unroll
    for (int c = tid; c < CHUNKS; c += THREADS) {
        int r = c / CHUNKS_PER_ROW, cc = (c % CHUNKS_PER_ROW) * 2;
        int rem = col_ext - cc;
        int bytes = (r < row_ext && rem > 0) ? (rem >= 2 ? 16 : 8) : 0;
        const double* g = bytes ? src + (size_t)r * ld + cc : src;
        cp_async16(sm + r * PITCH + cc, g, bytes);
    }
}

template <int BM, int BN, int BK, int WM, int WN, int STAGES, bool TA, bool TB>
__global__ void __launch_bounds__(GemmCfg<BM, BN, BK, WM, WN, STAGES, TA, TB>::THREADS)
gemm_f64_kernel(GemmArgs g) {
    using Cfg = GemmCfg<BM, BN, BK, WM, WN, STAGES, TA, TB>;
    extern __shared__ __align__(16) double smem[];
    // 1-D grid over the tiles. Tiles whose k range is longest are dispatched first (a triangular operand makes the work per
    // tile vary by the number of row / column tiles; heavy-first keeps the last wave short).
    const int tiles_m = (g.M + BM - 1) / BM, tiles_n = (g.N + BN - 1) / BN;
    int mt, nt;
    if (g.b_upper || g.b_lower) {  // work varies with the column tile: column tiles are the slow index
        nt = blockIdx.x / tiles_m; mt = blockIdx.x - nt * tiles_m;
        if (g.b_upper) nt = tiles_n - 1 - nt;
    } else {
        // Row tiles are the slow index INSIDE a band of `band_n` column tiles: a wave of CTAs then covers every row tile of one band, whose slice of op(B)
        // (band_n * BN columns x K, sized for the L2 by gemm_f64) is read from DRAM once instead of once per wave (L*Z at C3 with all 79 column tiles under
        // every row tile: 305 MB of DRAM traffic for 176 MB of operands, profiles/r01_ncu_gemm64.txt). op(A)'s row blocks are re-read per band from the L2.
        const int bw = (g.band_n > 0 && g.band_n < tiles_n) ? g.band_n : tiles_n;
        const int per_band = tiles_m * bw;
        const int band = blockIdx.x / per_band, r = blockIdx.x - band * per_band;
        const int w = min(bw, tiles_n - band * bw);  // the last band may be narrower
        mt = r / w; nt = band * bw + (r - mt * w);
        if (g.a_lower) mt = tiles_m - 1 - mt;
    }
    const int m0 = mt * BM, n0 = nt * BN;
    if (g.batch > 1) {
        const int z = blockIdx.z;
        g.A += z * g.strideA; g.B += z * g.strideB; g.C += z * g.strideC;
        if (z == g.batch - 1) {
            if (g.M_last > 0) g.M = g.M_last;
            if (g.N_last > 0) g.N = g.N_last;
            if (g.K_last > 0) g.K = g.K_last;
        }
        if (m0 >= g.M || n0 >= g.N) return;
    }
    if (g.lower_only && n0 >= m0 + BM) return;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm0 = (warp % Cfg::WARPS_M) * WM, wn0 = (warp / Cfg::WARPS_M) * WN;
    const int grp = lane >> 2, tig = lane & 3;

    int k_end = g.K, k_begin = 0;
    if (g.a_lower) k_end = min(k_end, g.m_off + m0 + BM);
    if (g.b_upper) k_end = min(k_end, g.n_off + n0 + BN);
    if (g.a_upper) k_begin = max(k_begin, g.m_off + m0);
    if (g.b_lower) k_begin = max(k_begin, g.n_off + n0);
    k_begin = (k_begin / BK) * BK;
    const int KT = k_end > k_begin ? (k_end - k_begin + BK - 1) / BK : 0;

    double* As = smem;
    double* Bs = smem + STAGES * Cfg::A_ELEMS;

    auto issue = [&](int kt, int stage) {
        const int k0 = k_begin + kt * BK;
        double* a = As + stage * Cfg::A_ELEMS;
        double* b = Bs + stage * Cfg::B_ELEMS;
        if (TA)  // op(A)[m][k] = A[k + m*lda]: rows are m, contiguous k
            load_tile<BM, BK, Cfg::A_PITCH, Cfg::THREADS>(a, g.A + (size_t)m0 * g.lda + k0, g.lda, g.M - m0, k_end - k0, tid);
        else     // A[m + k*lda]: rows are k, contiguous m
            load_tile<BK, BM, Cfg::A_PITCH, Cfg::THREADS>(a, g.A + (size_t)k0 * g.lda + m0, g.lda, k_end - k0, g.M - m0, tid);
        if (TB)  // op(B)[k][n] = B[n + k*ldb]: rows are k, contiguous n
            load_tile<BK, BN, Cfg::B_PITCH, Cfg::THREADS>(b, g.B + (size_t)k0 * g.ldb + n0, g.ldb, k_end - k0, g.N - n0, tid);
        else     // B[k + n*ldb]: rows are n, contiguous k
            load_tile<BN, BK, Cfg::B_PITCH, Cfg::THREADS>(b, g.B + (size_t)n0 * g.ldb + k0, g.ldb, g.N - n0, k_end - k0, tid);
    };

    constexpr int MI = WM / 8, NI = WN / 8;
    double acc[MI][NI][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NI; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < KT) issue(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < KT; ++kt) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        const int nxt = kt + STAGES - 1;
        if (nxt < KT) issue(nxt, nxt % STAGES);
        cp_async_commit();
        const double* a = As + (kt % STAGES) * Cfg::A_ELEMS;
        const double* b = Bs + (kt % STAGES) * Cfg::B_ELEMS;
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double af[MI], bf[NI];
#pragma unroll
            for (int i = 0; i < MI; ++i) {
                int m = wm0 + i * 8 + grp;
                af[i] = TA ? a[m * Cfg::A_PITCH + kk + tig] : a[(kk + tig) * Cfg::A_PITCH + m];
            }
#pragma unroll
            for (int j = 0; j < NI; ++j) {
                int n = wn0 + j * 8 + grp;
                bf[j] = TB ? b[(kk + tig) * Cfg::B_PITCH + n] : b[n * Cfg::B_PITCH + kk + tig];
            }
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NI; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
    }
    cp_async_wait<0>();

    // epilogue: thread holds rows grp, columns 2*tig, 2*tig+1 of each 8x8 tile. With beta != 0 the C values of half a warp tile are
    // loaded in one batch before anything is stored: load -> fma -> store per element would serialise 2*MI*NI memory round trips
    // (the compiler must assume the store aliases the next load), which is the whole cost of a short-K update (19 us at K = 64).
    constexpr int MH = (MI + 1) / 2;
#pragma unroll
    for (int ih = 0; ih < MI; ih += MH) {
        double cv[MH][NI][2];
        if (g.beta != 0.0) {
#pragma unroll
            for (int ii = 0; ii < MH; ++ii) {
                const int row = m0 + wm0 + (ih + ii) * 8 + grp;
#pragma unroll
                for (int j = 0; j < NI; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int col = n0 + wn0 + j * 8 + 2 * tig + e;
                        cv[ii][j][e] = (ih + ii < MI && row < g.M && col < g.N) ? __ldcg(g.C + (size_t)col * g.ldc + row) : 0.0;
                    }
            }
        }
#pragma unroll
        for (int ii = 0; ii < MH; ++ii) {
            if (ih + ii >= MI) continue;
            const int row = m0 + wm0 + (ih + ii) * 8 + grp;
            if (row >= g.M) continue;
#pragma unroll
            for (int j = 0; j < NI; ++j) {
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int col = n0 + wn0 + j * 8 + 2 * tig + e;
                    if (col >= g.N) continue;
                    double v = g.alpha * acc[ih + ii][j][e];
                    if (g.beta != 0.0) v += g.beta * cv[ii][j][e];
                    g.C[(size_t)col * g.ldc + row] = v;
                }
            }
        }
    }
}

// Host launcher: picks the tile shape from the output size.
cudaError_t gemm_f64(cudaStream_t st, bool ta, bool tb, const GemmArgs& g);

}  // namespace gpirt
