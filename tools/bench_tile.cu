// Where does a 64x64 tile solve spend its time? clock64-instrumented copy of k_trsm_left<false>.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int NB = 64;
__global__ void __launch_bounds__(256) k_tile(const double* __restrict__ L11, int ldl, int kb, double* __restrict__ B, int ldb, int nrhs, long long* t, int variant) {
    __shared__ double Ls[NB][NB + 1];
    __shared__ double vec[2][NB];
    __shared__ double inv[NB];
    long long c0t = clock64();
    const int tr = threadIdx.x & 15, tc = threadIdx.x >> 4;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = tr + 16 * p, c = tc + 16 * q;
            Ls[r][c] = (r < kb && c < kb) ? (c <= r ? L11[(size_t)c * ldl + r] : 0.0) : (r == c ? 1.0 : 0.0);
        }
    const int c0 = blockIdx.x * 64;
    double a[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = tr + 16 * p, c = c0 + tc + 16 * q;
            a[p][q] = (r < kb && c < nrhs) ? B[(size_t)c * ldb + r] : 0.0;
        }
    __syncthreads();
    if (threadIdx.x < NB) inv[threadIdx.x] = 1.0 / Ls[threadIdx.x][threadIdx.x];
    __syncthreads();
    long long c1t = clock64();
#pragma unroll
    for (int kq = 0; kq < 4; ++kq) {
        for (int kk = 0; kk < 16; ++kk) {
            const int k = 16 * kq + kk;
            double* xrow = vec[k & 1];
            if (tr == kk) {
                if (variant == 0) {
                    const double d = Ls[k][k];
#pragma unroll
                    for (int q = 0; q < 4; ++q) { const double v = a[kq][q] / d; a[kq][q] = v; xrow[tc + 16 * q] = v; }
                } else {
                    const double d = inv[k];
#pragma unroll
                    for (int q = 0; q < 4; ++q) { const double v = a[kq][q] * d; a[kq][q] = v; xrow[tc + 16 * q] = v; }
                }
            }
            __syncthreads();
            double xc[4], lr[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) { xc[p] = xrow[tc + 16 * p]; lr[p] = Ls[tr + 16 * p][k]; }
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const int r = tr + 16 * p;
                if (variant == 2 || r > k) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) a[p][q] -= lr[p] * xc[q];
                }
            }
        }
    }
    long long c2t = clock64();
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int r = tr + 16 * p, c = c0 + tc + 16 * q;
            if (r < kb && c < nrhs) B[(size_t)c * ldb + r] = a[p][q];
        }
    long long c3t = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { t[0] = c1t - c0t; t[1] = c2t - c1t; t[2] = c3t - c2t; }
}
int main() {
    const int n = 2000, N = 1001, ld = 2000;
    double *L, *B; long long* t;
    cudaMalloc(&L, (size_t)ld * n * 8); cudaMalloc(&B, (size_t)ld * N * 8); cudaMalloc(&t, 64);
    cudaMemset(L, 0, (size_t)ld * n * 8); cudaMemset(B, 0, (size_t)ld * N * 8);
    for (int variant = 0; variant < 3; ++variant) {
        for (int rep = 0; rep < 3; ++rep) {
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            cudaEventRecord(e0);
            k_tile<<<16, 256>>>(L + 64 + 64 * (size_t)ld, ld, 64, B + 64, ld, N, t, variant);
            cudaEventRecord(e1); cudaDeviceSynchronize();
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            long long h[3]; cudaMemcpy(h, t, 24, cudaMemcpyDeviceToHost);
            printf("variant %d rep %d: %.2f us | cycles load %lld loop %lld (%.0f/step) store %lld\n", variant, rep, ms * 1e3, h[0], h[1], h[1] / 64.0, h[2]);
        }
    }
    return 0;
}
