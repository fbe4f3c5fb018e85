// Tile-shape sweep of the fp64 DMMA GEMM on the three big products of a C3 sweep. Links ../gpirt_b200/build/linalg.o.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/bench_gemm tools/bench_gemm.cu gpirt_b200/build/linalg.o
#include <cstdio>
#include <vector>
#include "../gpirt_b200/csrc/gemm.cuh"
using namespace gpirt;
template <int BM, int BN, int BK, int WM, int WN, int ST, bool TA, bool TB>
float run(const GemmArgs& g, int reps) {
    using Cfg = GemmCfg<BM, BN, BK, WM, WN, ST, TA, TB>;
    auto kern = gemm_f64_kernel<BM, BN, BK, WM, WN, ST, TA, TB>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES) != cudaSuccess) return -1.f;
    dim3 grid(((g.N + BN - 1) / BN) * ((g.M + BM - 1) / BM), 1, 1);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES>>>(g); if (cudaDeviceSynchronize() != cudaSuccess) return -2.f;
    cudaEventRecord(e0); for (int r = 0; r < reps; ++r) kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES>>>(g); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms / reps;
}
#define ROW(BM, BN, BK, WM, WN, ST)                                                                                         \
    {                                                                                                                        \
        float a = run<BM, BN, BK, WM, WN, ST, false, false>(lz, 10), b = run<BM, BN, BK, WM, WN, ST, true, false>(ftw, 10),  \
              c = run<BM, BN, BK, WM, WN, ST, true, true>(zlp, 10);                                                          \
        printf("%3dx%3dx%2d warp %2dx%2d st%d | LZ %.3f ms %5.1f TF | FtW %.3f ms %5.1f TF | ZLpost %.3f ms %5.1f TF\n", BM, BN, BK, WM, WN, ST, a, \
               1.0 * n * n * m / a / 1e9, b, 2.0 * N * n * m / b / 1e9, c, 1.0 * N * N * m / c / 1e9);                       \
    }
int main() {
    const int n = 2000, N = 1001, m = 5000, ld = 2000, ldN = 1008, ldm = 5008;
    double *L, *Z, *F, *W, *O, *Lp;
    cudaMalloc(&L, (size_t)ld * n * 8); cudaMalloc(&Z, (size_t)ld * m * 8); cudaMalloc(&F, (size_t)ld * m * 8);
    cudaMalloc(&W, (size_t)ld * N * 8); cudaMalloc(&O, (size_t)ldm * 2048 * 8); cudaMalloc(&Lp, (size_t)ldN * N * 8);
    cudaMemset(L, 0, (size_t)ld * n * 8); cudaMemset(Z, 0, (size_t)ld * m * 8); cudaMemset(F, 0, (size_t)ld * m * 8);
    cudaMemset(W, 0, (size_t)ld * N * 8); cudaMemset(Lp, 0, (size_t)ldN * N * 8);
    GemmArgs lz{}; lz.M = n; lz.N = m; lz.K = n; lz.A = L; lz.lda = ld; lz.B = Z; lz.ldb = ld; lz.C = F; lz.ldc = ld; lz.alpha = 1; lz.a_lower = 1;
    GemmArgs ftw{}; ftw.M = m; ftw.N = N; ftw.K = n; ftw.A = F; ftw.lda = ld; ftw.B = W; ftw.ldb = ld; ftw.C = O; ftw.ldc = ldm; ftw.alpha = 1;
    GemmArgs zlp{}; zlp.M = m; zlp.N = N; zlp.K = N; zlp.A = Z; zlp.lda = ld; zlp.B = Lp; zlp.ldb = ldN; zlp.C = O; zlp.ldc = ldm; zlp.alpha = 1; zlp.beta = 1; zlp.b_upper = 1;
    ROW(128, 128, 16, 64, 32, 3)
    ROW(64, 64, 16, 32, 32, 3)
    ROW(64, 64, 16, 32, 32, 4)
    ROW(64, 64, 32, 32, 32, 3)
    ROW(128, 64, 16, 32, 32, 3)
    ROW(64, 128, 16, 32, 32, 3)
    ROW(128, 64, 16, 32, 32, 4)
    ROW(128, 64, 32, 32, 32, 3)
    ROW(128, 128, 16, 32, 32, 3)
    ROW(128, 128, 16, 32, 64, 3)
    ROW(64, 64, 16, 32, 16, 3)
    ROW(96, 64, 16, 32, 32, 3)
    return 0;
}
