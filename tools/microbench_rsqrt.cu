// Dependent-chain latency (cycles) of the pivot step of the tile Cholesky: rsqrt(double) as CUDA provides it vs an fp32-seeded
// coupled sqrt / rsqrt (Goldschmidt) iteration; plus accuracy of the latter against sqrt() / 1.0/sqrt().
#include <cmath>
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void sqrt_rsqrt_lib(double d, double& s, double& r) {
    r = rsqrt(d);
    s = d * r;
    s = fma(fma(-s, s, d), 0.5 * r, s);
}
__device__ __forceinline__ void sqrt_rsqrt_gs(double d, double& s, double& r) {
    float rf;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rf) : "f"((float)d));
    const double r0 = (double)rf;
    double g = d * r0, h = 0.5 * r0;
    double e = fma(-g, h, 0.5);
    g = fma(g, e, g); h = fma(h, e, h);
    e = fma(-g, h, 0.5);
    g = fma(g, e, g); h = fma(h, e, h);
    // one correction of g against d (as the library version does), r = 2h
    s = fma(fma(-g, g, d), h, g);
    r = h + h;
}
template <int V>
__global__ void k_chain(double* out, long long* cyc, int iters, double seed) {
    double x = seed + threadIdx.x * 1e-3, acc = 0.0;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        double s, r;
        if (V == 0) sqrt_rsqrt_lib(x, s, r); else sqrt_rsqrt_gs(x, s, r);
        acc += s;
        x = fma(r, 3.0, 1.5);  // next pivot depends on this one's reciprocal root
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[V] = t1 - t0;
    out[threadIdx.x + 32 * V] = x + acc;
}
__global__ void k_acc(const double* d, double* s, double* r, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) sqrt_rsqrt_gs(d[i], s[i], r[i]);
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, 64 * 8); cudaMalloc(&cyc, 16);
    const int iters = 4000;
    for (int rep = 0; rep < 2; ++rep) { k_chain<0><<<1, 32>>>(out, cyc, iters, 1.1); k_chain<1><<<1, 32>>>(out, cyc, iters, 1.1); cudaDeviceSynchronize(); }
    long long h[2]; cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
    printf("pivot chain cycles per step: library rsqrt %.1f, fp32-seeded Goldschmidt %.1f\n", h[0] / (double)iters, h[1] / (double)iters);
    const int n = 1 << 20;
    double *hd = new double[n], *hs = new double[n], *hr = new double[n], *d, *s, *r;
    unsigned long long st = 88172645463325252ull;
    for (int i = 0; i < n; ++i) { st ^= st << 13; st ^= st >> 7; st ^= st << 17; hd[i] = exp(((st >> 11) * (1.0 / 9007199254740992.0) - 0.5) * 80.0); }
    cudaMalloc(&d, n * 8); cudaMalloc(&s, n * 8); cudaMalloc(&r, n * 8);
    cudaMemcpy(d, hd, n * 8, cudaMemcpyHostToDevice);
    k_acc<<<n / 256, 256>>>(d, s, r, n); cudaMemcpy(hs, s, n * 8, cudaMemcpyDeviceToHost); cudaMemcpy(hr, r, n * 8, cudaMemcpyDeviceToHost);
    double es = 0, er = 0; int exact = 0;
    for (int i = 0; i < n; ++i) {
        double ts = sqrt(hd[i]), tr = (double)(1.0L / sqrtl((long double)hd[i]));
        es = fmax(es, fabs(hs[i] - ts) / ts); er = fmax(er, fabs(hr[i] - tr) / tr); exact += hs[i] == ts;
    }
    printf("accuracy over %d values in [e^-40, e^40]: sqrt max rel err %.3g (%.2f %% correctly rounded), rsqrt max rel err %.3g\n", n, es, 100.0 * exact / n, er);
    return 0;
}
