// Microbenchmarks that size the FP64 roofline on B200: DFMA issue rate, DMMA.8x8x4 rate, and the cost
// of the libm calls the probit likelihood leans on (erfc, log, exp). Build: see tools/run_microbench.sh
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)

__global__ void k_dfma(double* out, int iters, double a, double b) {
    double c[8];
    for (int k = 0; k < 8; ++k) c[k] = threadIdx.x + k;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) c[k] = fma(c[k], a, b);
    }
    double s = 0; for (int k = 0; k < 8; ++k) s += c[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma(double* out, int iters, double a, double b) {
    double c[8][2];
    for (int k = 0; k < 8; ++k) { c[k][0] = threadIdx.x; c[k][1] = k; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[k][0]), "+d"(c[k][1]) : "d"(a), "d"(b));
    }
    double s = 0; for (int k = 0; k < 8; ++k) s += c[k][0] + c[k][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP>
__global__ void k_fn(double* out, int iters, double x0) {
    double x[4]; for (int k = 0; k < 4; ++k) x[k] = x0 + 0.001 * (threadIdx.x + 32 * k);
    double acc = 0;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            double v;
            if (OP == 0) v = erfc(x[k]);
            else if (OP == 1) v = log(x[k] + 2.5);
            else if (OP == 2) v = exp(-x[k]);
            else if (OP == 3) v = normcdf(x[k]);
            else { double s, c; sincos(x[k], &s, &c); v = s + c; }
            acc += v; x[k] += 1e-7 * v;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    printf("device %s SMs %d clock %d kHz\n", p.name, sms, p.clockRate);
    double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 1024));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    for (int threads : {256, 512, 1024}) {
        int blocks = sms * (2048 / threads);
        int iters = 20000;
        k_dfma<<<blocks, threads>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
        cudaEventRecord(e0); k_dfma<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
        cudaEventElapsedTime(&ms, e0, e1);
        printf("DFMA  threads/blk %4d: %.2f TFLOP/s\n", threads, 2.0 * 8 * iters * (double)blocks * threads / ms / 1e9);
        k_dmma<<<blocks, threads>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
        cudaEventRecord(e0); k_dmma<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
        cudaEventElapsedTime(&ms, e0, e1);
        printf("DMMA884 threads/blk %4d: %.2f TFLOP/s\n", threads, 2.0 * 256 * 8 * iters * (double)blocks * (threads / 32) / ms / 1e9);
    }
    const char* names[5] = {"erfc", "log", "exp", "normcdf", "sincos"};
    for (int op = 0; op < 5; ++op) {
        int threads = 256, blocks = sms * 8, iters = 2000;
        auto launch = [&](int it) {
            switch (op) { case 0: k_fn<0><<<blocks, threads>>>(out, it, 0.3); break; case 1: k_fn<1><<<blocks, threads>>>(out, it, 0.3); break;
                          case 2: k_fn<2><<<blocks, threads>>>(out, it, 0.3); break; case 3: k_fn<3><<<blocks, threads>>>(out, it, 0.3); break;
                          default: k_fn<4><<<blocks, threads>>>(out, it, 0.3); }
        };
        launch(10); CK(cudaDeviceSynchronize());
        cudaEventRecord(e0); launch(iters); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
        cudaEventElapsedTime(&ms, e0, e1);
        double calls = 4.0 * iters * (double)blocks * threads;
        printf("%-8s %.1f Gcalls/s  (%.1f SM-cycles per warp-call at %d MHz)\n", names[op], calls / ms / 1e6,
               (ms * 1e-3 * p.clockRate * 1e3 * sms) / (calls / 32), p.clockRate / 1000);
    }
    return 0;
}
