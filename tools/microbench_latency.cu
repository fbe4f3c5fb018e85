// Dependent-chain latencies on B200 (cycles): DFMA, DDIV, DSQRT, smem round trip, __syncthreads — what bounds the
// 64-step elimination chains of the panel kernels and the per-cell chains of the ESS kernels.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_lat(double* out, long long* cyc, int iters, double seed) {
    __shared__ double sm[256];
    double x = seed + threadIdx.x * 1e-9;
    long long t0, t1;
    // DFMA chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = fma(x, 1.0000001, 1e-9);
    t1 = clock64(); if (threadIdx.x == 0) cyc[0] = t1 - t0;
    // DDIV chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = 1.5 / x + 1.0;
    t1 = clock64(); if (threadIdx.x == 0) cyc[1] = t1 - t0;
    // DSQRT chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = sqrt(x + 2.0);
    t1 = clock64(); if (threadIdx.x == 0) cyc[2] = t1 - t0;
    // smem store -> sync -> load chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { sm[threadIdx.x] = x; __syncthreads(); x = sm[(threadIdx.x + 1) & 255] + 1.0; __syncthreads(); }
    t1 = clock64(); if (threadIdx.x == 0) cyc[3] = t1 - t0;
    // sync only
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { __syncthreads(); }
    t1 = clock64(); if (threadIdx.x == 0) cyc[4] = t1 - t0;
    // drcp chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = __drcp_rn(x) + 1.0;
    t1 = clock64(); if (threadIdx.x == 0) cyc[5] = t1 - t0;
    // DADD chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = x + 1.0000001;
    t1 = clock64(); if (threadIdx.x == 0) cyc[6] = t1 - t0;
    // FFMA chain (fp32) for comparison
    float y = (float)x; t0 = clock64();
    for (int i = 0; i < iters; ++i) y = fmaf(y, 1.0000001f, 1e-9f);
    t1 = clock64(); if (threadIdx.x == 0) cyc[7] = t1 - t0;
    out[threadIdx.x] = x + y;
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, 256 * 8); cudaMalloc(&cyc, 8 * 8);
    const int iters = 2000;
    for (int threads : {32, 256}) {
        k_lat<<<1, threads>>>(out, cyc, iters, 1.1); cudaDeviceSynchronize();
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0); k_lat<<<1, threads>>>(out, cyc, iters, 1.1); cudaEventRecord(e1); cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long h[8]; cudaMemcpy(h, cyc, 64, cudaMemcpyDeviceToHost);
        long long tot = 0; for (int k = 0; k < 8; ++k) tot += h[k];
        printf("threads %3d: per-iteration cycles  DFMA %.1f  DDIV(+add) %.1f  DSQRT(+add) %.1f  smem+2sync %.1f  sync %.1f  DRCP(+add) %.1f  DADD %.1f  FFMA %.1f | kernel %.3f ms, %lld cycles -> %.0f MHz\n",
               threads, h[0] / (double)iters, h[1] / (double)iters, h[2] / (double)iters, h[3] / (double)iters, h[4] / (double)iters,
               h[5] / (double)iters, h[6] / (double)iters, h[7] / (double)iters, ms, tot, tot / (ms * 1e3));
    }
    return 0;
}
