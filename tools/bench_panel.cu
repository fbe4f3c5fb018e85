// Phase clocks of one Cholesky panel launch (CTA 1): compiled with the kernels' translation unit and -DGPIRT_PANEL_CLOCKS.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DGPIRT_PANEL_CLOCKS -o tools/bench_panel tools/bench_panel.cu
#include "../gpirt_b200/csrc/linalg.cu"

#include <vector>
using namespace gpirt;

int main() {
    const int n = 2000, ld = 2000;
    std::vector<double> th(n), hA((size_t)ld * n);
    for (int i = 0; i < n; ++i) th[i] = -3.0 + 6.0 * ((i * 7919) % n) / n;
    double *x, *A; int *info, *ws;
    cudaMalloc(&x, n * 8); cudaMalloc(&A, (size_t)ld * n * 8); cudaMalloc(&info, 4); cudaMalloc(&ws, 4 * chol_ws_ints(n));
    cudaMemcpy(x, th.data(), n * 8, cudaMemcpyHostToDevice);
    double sds[3] = {3, 3, 3};
    for (int rep = 0; rep < 3; ++rep) {
        build_K(0, x, n, x, n, sds, A, ld, 1e-6);
        cudaMemset(info, 0, 4); cudaMemset(ws, 0, 4 * chol_ws_ints(n));
        cudaDeviceSynchronize();
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        unsigned long long z4[4] = {0, 0, 0, 0}, pc[4];
        cudaMemcpyToSymbol(g_potf2_clk, z4, sizeof(z4));
        cudaEventRecord(e0);
        k_chol_panel<<<32, TILE_THREADS>>>(A, ld, 64, n - 64, info, 0, ws);
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long c[8];
        cudaMemcpyFromSymbol(c, g_panel_clk, sizeof(c));
        cudaMemcpyFromSymbol(pc, g_potf2_clk, sizeof(pc));
        printf("potf2 per step: diag factor %.0f, owner solve %.0f, update (thread 15,15) %.0f cycles\n", pc[0] / 16.0, pc[1] / 15.0, pc[2] / 15.0);
        printf("panel launch %.1f us; cycles: load %lld potf2 %lld stage %lld solve %lld store %lld total %lld\n", ms * 1e3, c[1] - c[0], c[2] - c[1],
               c[3] - c[2], c[4] - c[3], c[5] - c[4], c[5] - c[0]);
    }
    return 0;
}
