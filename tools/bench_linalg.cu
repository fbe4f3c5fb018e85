// Warm, event-timed costs of the dense building blocks (chol, trsm, gemm) at the C3 sizes. Links ../gpirt_b200/build/linalg.o.
#include <cstdio>
#include <vector>
#include <cmath>
#include "../gpirt_b200/csrc/linalg.cuh"
using namespace gpirt;
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)
template <class F> float timeit(F f, int reps) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); for (int r = 0; r < reps; ++r) f(); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms / reps;
}
int main() {
    const int n = 2000, N = 1001, m = 5000, ld = pad_ld(n), ldN = pad_ld(N);
    std::vector<double> th(n); for (int i = 0; i < n; ++i) th[i] = -5.0 + 0.01 * ((i * 7919) % 1001);
    double *dth, *K, *A, *B, *C; int* info;
    CK(cudaMalloc(&dth, n * 8)); CK(cudaMalloc(&K, (size_t)ld * n * 8)); CK(cudaMalloc(&A, (size_t)ld * n * 8));
    CK(cudaMalloc(&B, (size_t)ld * m * 8)); CK(cudaMalloc(&C, (size_t)ld * m * 8)); CK(cudaMalloc(&info, 4)); int* ws; CK(cudaMalloc(&ws, 4 * chol_ws_ints(n)));
    CK(cudaMemcpy(dth, th.data(), n * 8, cudaMemcpyHostToDevice)); CK(cudaMemset(info, 0, 4)); CK(cudaMemset(B, 0, (size_t)ld * m * 8));
    double sds[3] = {3, 3, 3};
    CK(build_K(0, dth, n, dth, n, sds, K, ld, 1e-6));
    auto reset = [&] { cudaMemcpyAsync(A, K, (size_t)ld * n * 8, cudaMemcpyDeviceToDevice, 0); };
    float t_copy = timeit(reset, 20);
    float t_chol = timeit([&] { reset(); chol_lower_blocked(0, A, ld, n, info, ws); }, 20) - t_copy;
    printf("chol n=%d: %.3f ms (%.2f TFLOP/s)\n", n, t_chol, n * 1.0 * n * n / 3 / t_chol / 1e9);
    int h; CK(cudaMemcpy(&h, info, 4, cudaMemcpyDeviceToHost)); printf("info %d\n", h);
    float t_trsm = timeit([&] { trsm_left_lower(0, A, ld, n, B, ld, N); }, 20);
    printf("trsm_left_lower n=%d nrhs=%d: %.3f ms\n", n, N, t_trsm);
    float t_trsmt = timeit([&] { trsm_left_lower_t(0, A, ld, n, B, ld, N); }, 20);
    printf("trsm_left_lower_t: %.3f ms\n", t_trsmt);
    GemmArgs g{}; g.M = n; g.N = m; g.K = n; g.A = A; g.lda = ld; g.B = B; g.ldb = ld; g.C = C; g.ldc = ld; g.alpha = 1; g.beta = 0; g.a_lower = 1;
    float t_g = timeit([&] { gemm_f64(0, false, false, g); }, 10);
    printf("gemm L*Z (tri) %dx%dx%d: %.3f ms (%.2f TFLOP/s)\n", n, m, n, t_g, 1.0 * n * n * m / t_g / 1e9);
    g.a_lower = 0;
    t_g = timeit([&] { gemm_f64(0, false, false, g); }, 10);
    printf("gemm dense NN %dx%dx%d: %.3f ms (%.2f TFLOP/s)\n", n, m, n, t_g, 2.0 * n * n * m / t_g / 1e9);
    // one trailing update, K = 64
    GemmArgs u{}; u.M = 1936; u.N = 1936; u.K = 64; u.A = A + 64; u.lda = ld; u.B = A + 64; u.ldb = ld; u.C = A + 64 + (size_t)64 * ld; u.ldc = ld; u.alpha = -1; u.beta = 1; u.lower_only = 1;
    float t_u = timeit([&] { gemm_f64(0, false, true, u); }, 50);
    printf("trailing update 1936^2 x 64: %.4f ms\n", t_u);
    u.M = u.N = 960;
    t_u = timeit([&] { gemm_f64(0, false, true, u); }, 50);
    printf("trailing update 960^2 x 64: %.4f ms\n", t_u);
    GemmArgs v{}; v.M = 1936; v.N = N; v.K = 64; v.A = A + 64; v.lda = ld; v.B = B; v.ldb = ld; v.C = B + 64; v.ldc = ld; v.alpha = -1; v.beta = 1;
    float t_v = timeit([&] { gemm_f64(0, false, false, v); }, 50);
    printf("trsm rhs update 1936 x 1001 x 64: %.4f ms\n", t_v);
    // empty-ish kernel launch cost in a stream
    float t_l = timeit([&] { build_K(0, dth, 1, dth, 1, sds, C, ld, 0.0); }, 200);
    printf("tiny kernel launch-to-launch: %.4f ms\n", t_l);
    return 0;
}
