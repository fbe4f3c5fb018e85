// The elliptical-slice samplers of the sweep as sm_100a kernels.
//   item-parallel (one CTA of 1-16 warps per item, warp-shuffle reductions, int8 responses):
//     draw_f            src/draw-f.cpp:5-83            ess_sparse_threadsafe per item
//     draw_beta         src/draw-beta.cpp:4-101
//     draw_threshold    src/draw_threshold.cpp:99-169,206-228
//     ll_bar_sparse     src/log-likelihood.cpp:50-64   (per-item totals, src/gpirtMCMC.cpp:324-331)
//   respondent-parallel (one CTA per respondent over the grid-row-major f* table):
//     draw_theta        src/draw-theta.cpp:4-84,112-170
// The shrink-bracket loop runs entirely on chip: the only per-proposal traffic is the re-read of the
// item's f / nu / y columns, which stay L1/L2 resident between proposals.
#include <cstdlib>
#include <algorithm>
#include "samplers.cuh"

#include "probit.cuh"

namespace gpirt {

#define TWO_PI_R 6.283185307179586476925286766559  // Rmath M_2PI (src/draw-f.cpp:19)

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------------
// Z[i + j*ld] = i-th standard normal of unit (unit0 + j) on `stream` (Box-Muller pairs from Philox)
__global__ void k_gen_normals(double* __restrict__ out, int ld, int rows, int cols, RngKey key, uint32_t stream, uint32_t unit0) {
    const int pairs = (rows + 1) >> 1;
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    int j = blockIdx.y;
    if (b >= pairs || j >= cols) return;
    double z0, z1;
    philox_normal2(key, stream, unit0 + j, b, z0, z1);
    double* col = out + (size_t)j * ld;
    col[2 * b] = z0;
    if (2 * b + 1 < rows) col[2 * b + 1] = z1;
}

cudaError_t launch_gen_normals(cudaStream_t st, double* out, int ld, int rows, int cols, RngKey key, uint32_t stream, uint32_t unit0) {
    if (rows <= 0 || cols <= 0) return cudaSuccess;
    dim3 grid(((rows + 1) / 2 + 127) / 128, cols);
    k_gen_normals<<<grid, 128, 0, st>>>(out, ld, rows, cols, key, stream, unit0);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// One CTA of WARPS warps per item. MODE selects what the ellipse moves: the item's f column, its 3 mean
// coefficients, or its C-1 threshold deltas. The shrink loop is a serial chain of likelihood passes, so the
// time of the block is (slowest item's proposals) x (latency of one pass): a pass is spread over enough
// threads that each one owns only a handful of cells, and cells are processed two at a time for ILP.
// Reductions are fixed-order (lane-strided partials -> xor butterfly -> warp partials summed in warp order).
template <int WARPS>
__device__ __forceinline__ double block_sum(double v, double* s_part) {
    v = warp_sum(v);
    if (WARPS == 1) return v;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();  // previous readers of s_part are done
    if (lane == 0) s_part[warp] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) t += s_part[w];
    return t;
}

// two sums at once (same fixed order): a and b come back on every thread. The partials are double-buffered by `parity` (the caller flips it on
// every call, uniformly over the CTA): a reduction then needs ONE barrier — the buffer it writes was last read two reductions ago, and those
// readers have passed the barrier of the reduction in between. A block reduction measured ~14 % of an 8-cell evaluation pass with two barriers.
template <int WARPS>
__device__ __forceinline__ void block_sum2(double& a, double& b, double (*s_part2)[2][WARPS], int& parity) {
    a = warp_sum(a);
    b = warp_sum(b);
    if (WARPS == 1) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double (*buf)[WARPS] = s_part2[parity];
    parity ^= 1;
    if (lane == 0) { buf[0][warp] = a; buf[1][warp] = b; }
    __syncthreads();
    double ta = 0.0, tb = 0.0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) { ta += buf[0][w]; tb += buf[1][w]; }
    a = ta; b = tb;
}

template <int MODE, int WARPS>
__device__ __forceinline__ void ess_one_item(const ItemView& v, const int j, double* tau, double2* tau2, double* s_part, double (*s_part2)[2][WARPS],
                                             int& red_parity) {
    constexpr int T = WARPS * 32;
    const int tid = threadIdx.x;
    const int J = v.item0 + j;
    const int C = v.C;
    for (int c = tid; c <= C; c += T) tau[c] = v.thr[(size_t)J * (C + 1) + c];
    __syncthreads();
    // tau2[y] = (tau[y-1], tau[y]): the two thresholds of a cell in one 16-byte shared-memory load indexed by the response itself; entry 0 (a missing
    // response, evaluated and not added in the packed loop) repeats category 1. Rewritten by whoever rewrites tau, in front of the same barrier.
    auto fill_tau2 = [&]() {
        for (int c = 1; c <= C; ++c) tau2[c] = make_double2(tau[c - 1], tau[c]);
        tau2[0] = tau2[1];
    };
    if (tid == 0) fill_tau2();
    __syncthreads();

    const int8_t* __restrict__ yj = v.y_im + (size_t)j * v.ldy;
    double* __restrict__ fj = v.f + (size_t)j * v.ldn;
    const double* __restrict__ nuj = (MODE == IB_F) ? v.nu + (size_t)j * v.ldn : nullptr;
    const double* __restrict__ theta = v.theta;
    const double* __restrict__ mur = (v.mu_rows && MODE != IB_BETA) ? v.mu_rows + (size_t)j * v.ldn : nullptr;
    const int n = v.n;
    // Early rejection (screened mode): every cell term is log(p + 1e-6) with p <= 1, i.e. at most log(1 + 1e-6) < 1.000001e-6, so a partial sum
    // over the first cells plus that cap for the cells still to come is an upper bound of the proposal's log-likelihood. The first proposals of
    // a shrink loop land far out on the ellipse (prior sds of 3 move g by several units: terms near log(1e-6)), and are certainly below log_y
    // after an eighth of the cells; the evaluation then stops there. Three phases: cells [0, c1), [c1, c2), [c2, CPT) per thread. The decision
    // is the one the full fp64 evaluation takes (test: screened == unscreened bit for bit), only rejections can come early.
    const int CPT = (n + T - 1) / T;                       // cells per thread
    const int c1 = CPT >= 4 ? max(1, CPT / 8) : CPT;       // fewer than 4 cells per thread: one phase
    const int c2 = CPT >= 4 ? max(c1 + 1, (3 * CPT) / 8) : CPT;
    const double b0 = v.beta[3 * (size_t)J], b1 = v.beta[3 * (size_t)J + 1], b2 = (v.p_beta == 3) ? v.beta[3 * (size_t)J + 2] : 0.0;
    double pb0 = b0, pb1 = b1, pb2 = b2;

    // one cell's term at the point (ce, se) of the ellipse (current == true: the state itself)
    auto term = [&](int i, double ce, double se, bool current) -> double {
        const int y = yj[i];
        if (y == 0) return 0.0;
        double fi = fj[i];
        if (MODE == IB_F && !current) fi = __dadd_rn(__dmul_rn(fi, ce), __dmul_rn(nuj[i], se));
        const double g = __dadd_rn(fi, mur ? mur[i] : mu_of(theta[i], pb0, pb1, pb2));
        return cell_ll(g, tau[y - 1], tau[y], y == 1, y == C);
    };
    auto loglik = [&](double ce, double se, bool current) -> double {
        double acc0 = 0.0, acc1 = 0.0;
        int i = tid;
        for (; i + T < n; i += 2 * T) {
            acc0 += term(i, ce, se, current);
            acc1 += term(i + T, ce, se, current);
        }
        if (i < n) acc0 += term(i, ce, se, current);
        return block_sum<WARPS>(acc0 + acc1, s_part);
    };

    // fp32 screening pass (probit.cuh) over the thread's cells c_begin <= c < c_end (cell c of thread tid is respondent tid + c * T): the
    // log-likelihood terms in fp32 (summed in fp64) and the bound on their distance to the fp64 terms, reduced over the CTA
    auto partial32 = [&](int c_begin, int c_end, double ce, double se, double& L, double& E) {
        double l0 = 0.0, l1 = 0.0;
        float e0 = 0.f, e1 = 0.f;
        auto one = [&](int i, double& l, float& e) {
            const int y = yj[i];
            if (y == 0) return;
            double fi = fj[i];
            if (MODE == IB_F) fi = __dadd_rn(__dmul_rn(fi, ce), __dmul_rn(nuj[i], se));
            const double g = __dadd_rn(fi, mur ? mur[i] : mu_of(theta[i], pb0, pb1, pb2));
            float t, et;
            cell_ll32(g, tau[y - 1], tau[y], t, et);
            l += (double)t;
            e += et;
        };
        // cells i and i + T as one packed pair (probit.cuh cell_ll32x2: the same (t, e) per cell, one instruction for both). A missing cell is
        // evaluated as category 1 and not added, so the pair's instruction stream has no branch.
        auto two = [&](int i) {
            const int y0 = yj[i], y1 = yj[i + T];
            double f0 = fj[i], f1 = fj[i + T];
            if (MODE == IB_F) {
                f0 = __dadd_rn(__dmul_rn(f0, ce), __dmul_rn(nuj[i], se));
                f1 = __dadd_rn(__dmul_rn(f1, ce), __dmul_rn(nuj[i + T], se));
            }
            const double g0 = __dadd_rn(f0, mur ? mur[i] : mu_of(theta[i], pb0, pb1, pb2));
            const double g1 = __dadd_rn(f1, mur ? mur[i + T] : mu_of(theta[i + T], pb0, pb1, pb2));
            const double2 th0 = tau2[y0], th1 = tau2[y1];
            float t0, t1, et0, et1;
            cell_ll32x2(g0, th0.x, th0.y, g1, th1.x, th1.y, t0, et0, t1, et1);
            l0 += (double)(y0 > 0 ? t0 : 0.0f);  // adding +0 leaves the sums as they are
            l1 += (double)(y1 > 0 ? t1 : 0.0f);
            e0 += y0 > 0 ? et0 : 0.0f;
            e1 += y1 > 0 ? et1 : 0.0f;
        };
        const int i_end = min(n, c_end * T);
        int i = tid + c_begin * T;
        for (; i + T < i_end; i += 2 * T) two(i);
        if (i < i_end) one(i, l0, e0);
        L = l0 + l1;
        E = (double)(e0 + e1);
        block_sum2<WARPS>(L, E, s_part2, red_parity);
    };
    auto loglik32 = [&](double ce, double se, double& L, double& E) { partial32(0, CPT, ce, se, L, E); };

    // prior draw for the low-dimensional blocks
    double nu_small[MAX_C];  // beta: 3 used; thresholds: C-1 used
    double delta[MAX_C];
    if (MODE == IB_BETA) {
        // items without observations keep beta and draw nothing (src/draw-beta.cpp:73,97-99)
        int any = 0;
        for (int i = tid; i < n; i += T) any |= (yj[i] != 0);
        if (!__syncthreads_or(any)) {
            if (tid == 0) { v.evals[J] = 0; v.margin[J] = CUDART_INF; }
            return;
        }
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const double z = v.z ? v.z[(size_t)J * v.z_stride + q] : philox_normal(v.key, ST_BETA_Z, J, q);
            const double sd = v.sds[3 * (size_t)J + q];
            nu_small[q] = __dmul_rn(sqrt(__dadd_rn(__dmul_rn(sd, sd), 1e-6)), z);  // chol(diag(sd^2+1e-6)) z, :82-86
        }
    }
    if (MODE == IB_THR) {
        // threshold_to_delta (src/log-likelihood.cpp:78-88), identity prior (src/draw_threshold.cpp:110-114)
        delta[0] = tau[1];
        for (int q = 1; q < C - 1; ++q) delta[q] = log(__dsub_rn(tau[q + 1], tau[q]));
        for (int q = 0; q < C - 1; ++q) nu_small[q] = v.z ? v.z[(size_t)J * v.z_stride + q] : philox_normal(v.key, ST_THR_Z, J, q);
        // the reference evaluates the current state through delta_to_threshold(delta) (:118)
        __syncthreads();
        if (tid == 0) {
            tau[1] = delta[0];
            for (int q = 1; q < C - 1; ++q) tau[q + 1] = __dadd_rn(tau[q], exp(delta[q]));
            fill_tau2();
        }
        __syncthreads();
    }

    const uint32_t ustream = MODE == IB_F ? ST_F_U : (MODE == IB_BETA ? ST_BETA_U : ST_THR_U);
    int overrun = 0;
    uint32_t k = 0;
    const double u = v.us.get(v.key, ustream, J, k++, &overrun);
    const double log_u = log(u);
    // the point (ce, se) of the ellipse as the block's parameters: beta' (registers) or tau' = delta_to_threshold(delta') (shared)
    auto set_params = [&](double ce_, double se_) {
        if (MODE == IB_BETA) {
            pb0 = __dadd_rn(__dmul_rn(b0, ce_), __dmul_rn(nu_small[0], se_));
            pb1 = __dadd_rn(__dmul_rn(b1, ce_), __dmul_rn(nu_small[1], se_));
            pb2 = __dadd_rn(__dmul_rn(b2, ce_), __dmul_rn(nu_small[2], se_));
        }
        if (MODE == IB_THR) {
            // no barrier in front: every evaluation ends in a block reduction, whose barrier all threads pass only after their last read of tau
            if (tid == 0) {  // delta_to_threshold(delta'), src/log-likelihood.cpp:66-76
                tau[1] = __dadd_rn(__dmul_rn(delta[0], ce_), __dmul_rn(nu_small[0], se_));
                for (int q = 1; q < C - 1; ++q)
                    tau[q + 1] = __dadd_rn(tau[q], exp(__dadd_rn(__dmul_rn(delta[q], ce_), __dmul_rn(nu_small[q], se_))));
                fill_tau2();
            }
            __syncthreads();
        }
    };
    // Unscreened: log_y = ll(current) + log(u) in fp64, every proposal in fp64. Screened: the current state is evaluated in fp32 with its bound
    // like the proposals, and a proposal is decided when |S' - S0 - log u| exceeds both bounds; only an undecided proposal makes the kernel
    // compute the two fp64 values (the current state's once) — the decision is the fp64 decision either way.
    double S0 = 0.0, E0 = 0.0, log_y = 0.0;
    bool have_exact0 = false;
    int evals = 1, n_exact = 0;  // n_exact: evaluations that ran in fp64
    if (v.screen) loglik32(1.0, 0.0, S0, E0);
    else { log_y = __dadd_rn(loglik(1.0, 0.0, true), log_u); have_exact0 = true; n_exact = 1; }
    double eps_min = 0.0, eps_max = TWO_PI_R;
    double eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ustream, J, k++, &overrun)));
    eps_min = __dsub_rn(eps, TWO_PI_R);
    double margin = CUDART_INF;
    double ce = 1.0, se = 0.0;
    bool accepted = false;
    for (int it = 0; it < ESS_MAX_PROPOSALS; ++it) {
        ce = cos(eps);
        se = sin(eps);
        set_params(ce, se);
        ++evals;
        bool decided = false, accept = false;
        if (v.screen) {
            double L32 = 0.0, E32 = 0.0;
            int c_done = 0;
            if (v.early_exit && c1 < CPT) {
                // noise of the fp64 passes this stands in for: their sums are at most 13.9 n in magnitude (every term >= log(1e-6 - eps))
                const double noise = 1e-10 * (13.9 * (double)n + fabs(S0) + 1.0);
                for (int ph = 0; ph < 2 && !decided; ++ph) {
                    const int c_to = ph == 0 ? c1 : c2;
                    double Lp, Ep;
                    partial32(c_done, c_to, ce, se, Lp, Ep);
                    L32 += Lp; E32 += Ep; c_done = c_to;
                    const double cap = 1.000001e-6 * (double)max(0, n - c_to * T);  // the cells still to come add at most this
                    const double d = (L32 + cap) - (S0 + log_u);
                    const double Eb = (E32 + E0) * 1.00002 + noise;
                    if (d < -Eb) { decided = true; margin = fmin(margin, -d - Eb); }
                }
            }
            if (!decided) {
                double Lp, Ep;
                partial32(c_done, CPT, ce, se, Lp, Ep);
                L32 += Lp; E32 += Ep;
                const double d = L32 - (S0 + log_u);
                // fp32 rounding of the E accumulations (relative 1e-5 at most) and the fp64 passes' own summation noise
                const double Eb = (E32 + E0) * 1.00002 + 1e-10 * (fabs(L32) + fabs(S0) + 1.0);
                if (d > Eb) { decided = true; accept = true; }
                else if (d < -Eb) { decided = true; }
                if (decided) margin = fmin(margin, fabs(d) - Eb);  // a lower bound of the true margin
            }
        }
        if (!decided) {
            if (!have_exact0) {  // the current state in fp64: its own parameters for the pass, then the proposal's again
                if (MODE == IB_BETA) { pb0 = b0; pb1 = b1; pb2 = b2; }
                if (MODE == IB_THR) {
                    __syncthreads();
                    if (tid == 0) {
                        tau[1] = delta[0];
                        for (int q = 1; q < C - 1; ++q) tau[q + 1] = __dadd_rn(tau[q], exp(delta[q]));
                        fill_tau2();
                    }
                    __syncthreads();
                }
                log_y = __dadd_rn(loglik(1.0, 0.0, true), log_u);
                set_params(ce, se);
                have_exact0 = true;
                ++n_exact;
            }
            const double llp = loglik(ce, se, false);  // identical on every thread of the CTA
            margin = fmin(margin, fabs(llp - log_y));
            accept = llp > log_y;
            ++n_exact;
        }
        if (accept) { accepted = true; break; }
        if (eps < 0.0) eps_min = eps; else eps_max = eps;
        eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ustream, J, k++, &overrun)));
        if (overrun) break;
    }
    if (tid == 0) {
        v.evals[J] = evals;
        v.margin[J] = margin;
        if (v.exact_evals) v.exact_evals[J] = n_exact;
        if (overrun) atomicExch(&v.flags->overrun, 1);
        else if (!accepted) atomicExch(&v.flags->stalled, 1);
    }
    // commit the last proposal (the reference's `out = f_prime`): on a stall/overrun this is still a
    // point of the ellipse, and the raised flag turns the call into an error at the next sync
    if (MODE == IB_F) {
        for (int i = tid; i < n; i += T) fj[i] = __dadd_rn(__dmul_rn(fj[i], ce), __dmul_rn(nuj[i], se));
    } else if (MODE == IB_BETA) {
        if (tid == 0) {
            v.beta[3 * (size_t)J] = pb0;
            v.beta[3 * (size_t)J + 1] = pb1;
            v.beta[3 * (size_t)J + 2] = pb2;
        }
    } else {
        __syncthreads();
        for (int c = 1 + tid; c < C; c += T) v.thr[(size_t)J * (C + 1) + c] = tau[c];
    }
}

// One CTA per item when the grid covers the items; with a smaller grid (launch_ess_mode: beta / threshold blocks, which share the SMs with the
// theta-only chain on the side stream) every CTA walks the items with the grid's stride. An item's arithmetic does not depend on which CTA ran it.
// 1024 resident threads per SM at least, i.e. at most 64 registers: the packed two-cell loop otherwise takes 76-80 in the beta / threshold kernels and
// drops them from four 256-thread CTAs per SM to three
template <int MODE, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 32 / WARPS) k_ess_items(ItemView v) {
    __shared__ double tau[MAX_C + 1];
    __shared__ double2 tau2[MAX_C + 1];
    __shared__ double s_part[WARPS];
    __shared__ double s_part2[2][2][WARPS];
    int red_parity = 0;
    for (int j = blockIdx.x; j < v.m_local; j += gridDim.x) {
        ess_one_item<MODE, WARPS>(v, j, tau, tau2, s_part, s_part2, red_parity);
        __syncthreads();  // tau / s_part are reused by the next item
    }
}

// warps per item: each thread should own <= ~8 cells of a pass
static int item_warps(int n) { return n <= 256 ? 1 : (n <= 1024 ? 4 : (n <= 4096 ? 8 : 16)); }

int g_ess_threads_per_sm = -1;  // beta / threshold ESS: resident threads per SM the grid is sized for (0: one CTA per item); -1: environment / default

template <int MODE>
static void launch_ess_mode(cudaStream_t st, const ItemView& v) {
    if (g_ess_threads_per_sm < 0) {
        const char* e = getenv("GPIRT_ESS_THREADS_PER_SM");
        g_ess_threads_per_sm = e ? atoi(e) : 0;
    }
    const int warps = item_warps(v.n);
    int grid = v.m_local;
    // The beta and threshold blocks run while the theta-only chain of the next sweep (Cholesky panels, small GEMMs) is on the side stream: a grid
    // that leaves a share of every SM's thread slots free lets those latency-bound kernels start at once instead of waiting for an ESS CTA to retire.
    if (MODE != IB_F && g_ess_threads_per_sm > 0) grid = std::min(grid, 148 * std::max(1, g_ess_threads_per_sm / (warps * 32)));
    switch (warps) {
        case 1: k_ess_items<MODE, 1><<<grid, 32, 0, st>>>(v); break;
        case 4: k_ess_items<MODE, 4><<<grid, 128, 0, st>>>(v); break;
        case 8: k_ess_items<MODE, 8><<<grid, 256, 0, st>>>(v); break;
        default: k_ess_items<MODE, 16><<<grid, 512, 0, st>>>(v); break;
    }
    ++g_kernel_launches;
}

cudaError_t launch_ess_items(cudaStream_t st, int block, const ItemView& v) {
    if (v.m_local <= 0) return cudaSuccess;
    if (block == IB_F) launch_ess_mode<IB_F>(st, v);
    else if (block == IB_BETA) launch_ess_mode<IB_BETA>(st, v);
    else launch_ess_mode<IB_THR>(st, v);
    return cudaGetLastError();
}

// per-item ll_bar_sparse with the current state. Grid-stride over the items: the Phi table (probit.cuh, phi_tab) is staged in shared memory once per CTA.
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_item_ll(ItemView v) {
    constexpr int T = WARPS * 32;
    __shared__ PhiNode s_phi[PHI_TAB_N];
    __shared__ LogNode s_log[128];
    __shared__ double tau[MAX_C + 1];
    __shared__ double s_part[WARPS];
    const int tid = threadIdx.x;
    load_phi_table(s_phi);
    for (int q = tid; q < 128; q += T) s_log[q] = g_log_tab[q];
    constexpr int U = 4;  // cells in flight per thread: their loads are issued together (ncu on the one-cell loop: long_scoreboard 7.8 warps per issue)
    for (int j = blockIdx.x; j < v.m_local; j += gridDim.x) {
        const int J = v.item0 + j;
        __syncthreads();  // tables staged; the previous item's tau / s_part are no longer read
        for (int c = tid; c <= v.C; c += T) tau[c] = v.thr[(size_t)J * (v.C + 1) + c];
        __syncthreads();
        const int8_t* __restrict__ yj = v.y_im + (size_t)j * v.ldy;
        const double* __restrict__ fj = v.f + (size_t)j * v.ldn;
        const double* __restrict__ theta = v.theta;
        const double b0 = v.beta[3 * (size_t)J], b1 = v.beta[3 * (size_t)J + 1], b2 = (v.p_beta == 3) ? v.beta[3 * (size_t)J + 2] : 0.0;
        double acc = 0.0;  // cells tid, tid + T, tid + 2T, ... added in this order
        for (int i0 = tid; i0 < v.n; i0 += U * T) {
            int y[U];
            double f[U], th[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * T;
                const bool in = i < v.n;
                y[u] = in ? (int)yj[i] : 0;
                f[u] = in ? fj[i] : 0.0;
                th[u] = in ? theta[i] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int c = y[u] > 0 ? y[u] : 1;  // a missing cell is evaluated as category 1 and not added
                const double g = __dadd_rn(f[u], mu_of(th[u], b0, b1, b2));
                const double t = cell_ll_tab2(g, tau[c - 1], tau[c], s_phi, s_log);
                if (y[u] > 0) acc += t;
            }
        }
        acc = block_sum<WARPS>(acc, s_part);
        if (tid == 0) v.item_ll[J] = acc;
    }
}

template <int WARPS>
static cudaError_t launch_item_ll_w(cudaStream_t st, const ItemView& v) {
    // a grid-stride kernel: exactly the CTAs that are resident at once (more would run as a second wave behind CTAs that loop over all their items)
    int per_sm = 0, dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_item_ll<WARPS>, WARPS * 32, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    const int grid = std::min(v.m_local, sms * per_sm);
    k_item_ll<WARPS><<<grid, WARPS * 32, 0, st>>>(v);
    ++g_kernel_launches;
    return cudaGetLastError();
}

cudaError_t launch_item_ll(cudaStream_t st, const ItemView& v) {
    if (v.m_local <= 0) return cudaSuccess;
    switch (item_warps(v.n)) {
        case 1: return launch_item_ll_w<1>(st, v);
        case 4: return launch_item_ll_w<4>(st, v);
        case 8: return launch_item_ll_w<8>(st, v);
        default: return launch_item_ll_w<16>(st, v);
    }
}

}  // namespace gpirt

namespace gpirt {

// ------------------------------------------------------------------------------------------------
// draw_theta: one CTA per respondent. The f* table is grid-row-major, so the row a proposal lands on
// is one contiguous run of m doubles per shard chunk (the reference gathers it with stride N).
constexpr int THETA_THREADS = 256;
int g_theta_threads = 0;  // 0: by the number of local respondents; 256 / 1024: forced (tests: both give the same bits)

__global__ void __launch_bounds__(128) k_theta_table(ThetaView v, double* __restrict__ table) {
    __shared__ PhiNode s_phi[PHI_TAB_N];
    load_phi_table(s_phi);
    __syncthreads();
    const int J = blockIdx.x * 128 + threadIdx.x;
    if (J >= v.m_total) return;
    const int C = v.C;
    const int r = J / v.m_chunk, jl = J - r * v.m_chunk;
    const double b0 = v.beta[3 * (size_t)J], b1 = v.beta[3 * (size_t)J + 1], b2 = v.beta[3 * (size_t)J + 2];
    const double* tau = v.thr + (size_t)J * (C + 1);
    for (int idx = blockIdx.y; idx < NGRID; idx += gridDim.y) {  // the Phi table is staged once per CTA, then reused over the CTA's grid rows
        const double fs = v.fstar_all[(size_t)r * v.chunk_elems + (size_t)idx * v.ld_chunk + jl];
        const double g = __dadd_rn(fs, mu_of(grid_value(idx), b0, b1, b2));
        double* out = table + ((size_t)idx * v.m_total + J) * C;
        double p_lo = 0.0;
        for (int c = 1; c <= C; ++c) {
            const double p_hi = (c == C) ? 1.0 : phi_tab(tau[c] - g, s_phi);
            out[c - 1] = cell_from_probs(p_lo, p_hi);
            p_lo = p_hi;
        }
    }
}

cudaError_t launch_theta_table(cudaStream_t st, const ThetaView& v, double* table) {
    const int bx = (v.m_total + 127) / 128;
    const int by = std::max(1, std::min((int)NGRID, (148 * 16 + bx - 1) / bx));  // ~16 CTAs of 128 threads per SM in one wave
    dim3 grid(bx, by);
    k_theta_table<<<grid, 128, 0, st>>>(v, table);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// The same table in fp32 with a rigorous bound per term (probit.cuh, "fp32 screening"): (t, e) with |t - fp64 term| <= e, one 8-byte record
// per term (a 5-byte form, fp32 term + 8-bit code of the bound in a second array, was measured: the second gather makes the latency-bound
// theta kernel 0.39 -> 0.62 ms at C3). The C + 1 tails
// of one (grid point, item) are shared by its C cells, so a cell costs ~1.2 tails and one lg2 on the fp32 / MUFU pipes instead of 1.2 fp64
// Phi and one fp64 log: the theta step's ESS decisions are screened on this table and only undecided proposals are evaluated in fp64.
__global__ void __launch_bounds__(128) k_theta_table32(ThetaView v, float2* __restrict__ table32) {
    // the C records of the CTA's 128 items are one contiguous run of the table: staged in shared memory, written out coalesced
    __shared__ float2 s_out[128 * MAX_C];
    const int J0 = blockIdx.x * 128, J = J0 + threadIdx.x;
    const int idx = blockIdx.y;
    const int C = v.C;
    if (J < v.m_total) {
        const int r = J / v.m_chunk, jl = J - r * v.m_chunk;
        const double fs = v.fstar_all[(size_t)r * v.chunk_elems + (size_t)idx * v.ld_chunk + jl];
        const double mu = mu_of(grid_value(idx), v.beta[3 * (size_t)J], v.beta[3 * (size_t)J + 1], v.beta[3 * (size_t)J + 2]);
        const double g = __dadd_rn(fs, mu);
        const double* tau = v.thr + (size_t)J * (C + 1);
        float2* out = s_out + threadIdx.x * C;
        float a = (float)__dsub_rn(tau[0], g);  // -inf: clamps to Q(9.25)
        float u = fminf(fabsf(a), 9.25f), qu = q_tail32(u);
        for (int c = 1; c <= C; ++c) {
            const float b = (float)__dsub_rn(tau[c], g);
            const float w = fminf(fabsf(b), 9.25f), qw = q_tail32(w);
            float t, e;
            cell_ll32_core(a, b, u, w, qu, qw, t, e);
            out[c - 1] = make_float2(t, e);
            a = b; u = w; qu = qw;
        }
    }
    __syncthreads();
    const int count = min(128, v.m_total - J0) * C;
    float2* dst = table32 + ((size_t)idx * v.m_total + J0) * C;
    for (int q = threadIdx.x; q < count; q += 128) dst[q] = s_out[q];
}

cudaError_t launch_theta_table32(cudaStream_t st, const ThetaView& v, float2* table32) {
    dim3 grid((v.m_total + 127) / 128, NGRID);
    k_theta_table32<<<grid, 128, 0, st>>>(v, table32);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// T threads per respondent: 256 when there are respondents enough to fill the GPU, 1024 for a rank with few of them (item-sharded runs:
// n / world respondents against m * world items). Both sum an evaluation in the SAME order — 1024 virtual lanes (lane l takes items l,
// l + 1024, ...), lanes t, t+256, t+512, t+768 combined as (S_t + S_t+256) + (S_t+512 + S_t+768), then the 8-warp block sum — so a chain
// does not depend on which one ran (bit-identical for any world size).
template <int T>
__global__ void __launch_bounds__(T, T == 256 ? 4 : 1) k_ess_theta(ThetaView v) {
    static_assert(T == 256 || T == 1024, "virtual-lane scheme");
    __shared__ double s_part[8];
    __shared__ double s_lane[T == 1024 ? 1024 : 1];
    __shared__ double s_part2[2][2][T / 32];
    int red_parity = 0;
    const int il = blockIdx.x;
    if (il >= v.n_local) return;
    const int i = v.resp0 + il;
    const int tid = threadIdx.x;
    const int C = v.C, m_total = v.m_total;
    // the respondent's row of y is read by every one of its ~12 evaluations: staged in shared memory once (when it fits), so an evaluation
    // is one independent gather per item instead of a y load followed by a dependent gather
    extern __shared__ __align__(16) int8_t s_y[];
    const int8_t* yrow = v.y_rm + (size_t)il * v.ldm;
    if (v.y_in_smem) {
        const int4* src = reinterpret_cast<const int4*>(yrow);  // rows are 16-byte aligned (ldm is a multiple of 16)
        int4* dst = reinterpret_cast<int4*>(s_y);
        for (int q = tid; q < (m_total + 15) / 16; q += T) dst[q] = src[q];
        __syncthreads();
        yrow = s_y;
    }
    const double* __restrict__ table = v.table;

    auto term = [&](int J, int idx, double ts) -> double {  // one item's term of compute_ll_sparse, src/draw-theta.cpp:14-22
        const int y = yrow[J];
        if (y == 0) return 0.0;
        if (table) return table[((size_t)idx * m_total + J) * C + (y - 1)];
        const int r = J / v.m_chunk, jl = J - r * v.m_chunk;
        const double fs = v.fstar_all[(size_t)r * v.chunk_elems + (size_t)idx * v.ld_chunk + jl];
        const double mu = mu_of(ts, v.beta[3 * (size_t)J], v.beta[3 * (size_t)J + 1], v.beta[3 * (size_t)J + 2]);
        const double g = __dadd_rn(fs, mu);  // f* already holds mu*; the reference adds it again (:17)
        return cell_ll(g, v.thr[(size_t)J * (C + 1) + y - 1], v.thr[(size_t)J * (C + 1) + y], y == 1, y == C);
    };
    auto loglik = [&](double th) -> double {
        const int idx = grid_index(th);
        const double ts = grid_value(idx);
        if (T == 256) {
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            int J = tid;
            for (; J + 768 < m_total; J += 1024) {
                a0 += term(J, idx, ts);
                a1 += term(J + 256, idx, ts);
                a2 += term(J + 512, idx, ts);
                a3 += term(J + 768, idx, ts);
            }
            if (J < m_total) a0 += term(J, idx, ts);          // the ragged last group: every item still goes to its own lane
            if (J + 256 < m_total) a1 += term(J + 256, idx, ts);
            if (J + 512 < m_total) a2 += term(J + 512, idx, ts);
            return block_sum<8>((a0 + a1) + (a2 + a3), s_part);
        } else {
            double a = 0.0;
            for (int J = tid; J < m_total; J += 1024) a += term(J, idx, ts);
            __syncthreads();  // previous readers of s_lane / s_part are done
            s_lane[tid] = a;
            __syncthreads();
            if (tid < 256) {
                double w = warp_sum((s_lane[tid] + s_lane[tid + 256]) + (s_lane[tid + 512] + s_lane[tid + 768]));
                if ((tid & 31) == 0) s_part[tid >> 5] = w;
            }
            __syncthreads();
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += s_part[w];
            return t;
        }
    };

    // screening pass: the evaluation's fp32 terms and their bounds from the (t, e) table, summed in fp64 (any order: the bound covers it)
    const float2* __restrict__ table32 = v.table32;
    auto loglik32 = [&](double th, double& S, double& E) {
        const float2* __restrict__ row = table32 + (size_t)grid_index(th) * m_total * C;
        double s0 = 0.0, s1 = 0.0, e0 = 0.0, e1 = 0.0;
        auto one = [&](int J, double& s, double& e) {
            const int y = yrow[J];
            if (y == 0) return;
            const float2 te = row[(size_t)J * C + (y - 1)];
            s += (double)te.x;
            e += (double)te.y;
        };
        int J = tid;
        for (; J + 3 * T < m_total; J += 4 * T) { one(J, s0, e0); one(J + T, s1, e1); one(J + 2 * T, s0, e0); one(J + 3 * T, s1, e1); }
        for (; J < m_total; J += T) one(J, s0, e0);
        S = s0 + s1;
        E = e0 + e1;
        block_sum2<T / 32>(S, E, s_part2, red_parity);
    };

    double Li;
    if (v.regime == 0) {
        const double sd = v.theta_prior_sds[2 * (size_t)i];
        Li = sqrt(__dadd_rn(1.0, __dmul_rn(sd, sd)));  // chol(V + sd^2), src/draw-theta.cpp:158,194
    } else {
        Li = sqrt(__dadd_rn(__dmul_rn(v.theta_os, v.theta_os), 1e-6));  // L_time at H = 1, src/cholesky-cache.cpp:29-33
    }
    const double z = v.z ? v.z[(size_t)i * v.z_stride] : philox_normal(v.key, ST_THETA_Z, i, 0);
    const double nu = __dmul_rn(Li, z);
    const double th0 = v.theta[i];
    int overrun = 0;
    uint32_t k = 0;
    const double u = v.us.get(v.key, ST_THETA_U, i, k++, &overrun);
    const double log_u = log(u);
    // Exact path: log_y = log(u) + ll(theta) in fp64, every proposal in fp64. Screened path (table32): the current state and the proposal
    // are both read from the fp32 table; the decision is taken there when |S' - S0 - log u| exceeds the two accumulated bounds, otherwise the
    // fp64 values are computed (the current state's once) and decide — so every decision is the fp64 decision.
    const bool screen = table32 != nullptr;
    double S0 = 0.0, E0 = 0.0, log_y = 0.0;
    bool have_exact0 = false;
    int n_exact = 0;
    if (screen) loglik32(th0, S0, E0);
    else { log_y = __dadd_rn(log_u, loglik(th0)); have_exact0 = true; n_exact = 1; }
    double eps_min = 0.0, eps_max = TWO_PI_R;
    double eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ST_THETA_U, i, k++, &overrun)));
    eps_min = __dsub_rn(eps, TWO_PI_R);
    int evals = 1;
    double margin = CUDART_INF, raw = th0;
    bool accepted = false;
    for (int it = 0; it < ESS_MAX_PROPOSALS; ++it) {
        raw = __dadd_rn(__dmul_rn(th0, cos(eps)), __dmul_rn(nu, sin(eps)));
        raw = fmax(-5.0, fmin(5.0, raw));
        ++evals;
        bool decided = false, accept = false;
        if (screen) {
            double Sp, Ep;
            loglik32(raw, Sp, Ep);
            const double d = Sp - (S0 + log_u);
            // + the fp64 pass's own summation / rounding noise (1e-16 relative per operation on sums of magnitude |S|)
            const double Eb = (Ep + E0) * 1.000001 + 1e-10 * (fabs(Sp) + fabs(S0) + 1.0);
            if (d > Eb) { decided = true; accept = true; }
            else if (d < -Eb) { decided = true; }
            if (decided) margin = fmin(margin, fabs(d) - Eb);  // a lower bound of the true margin
        }
        if (!decided) {
            if (!have_exact0) { log_y = __dadd_rn(log_u, loglik(th0)); have_exact0 = true; ++n_exact; }
            const double llp = loglik(raw);
            ++n_exact;
            margin = fmin(margin, fabs(llp - log_y));
            accept = llp > log_y;
        }
        if (accept) { accepted = true; break; }
        if (eps < 0.0) eps_min = eps; else eps_max = eps;
        eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ST_THETA_U, i, k++, &overrun)));
        if (overrun) break;
    }
    if (tid == 0) {
        v.theta_new[i] = grid_value(grid_index(raw));  // snapped to the grid, src/draw-theta.cpp:165-169
        v.evals[i] = evals;
        v.margin[i] = margin;
        if (v.exact_evals) v.exact_evals[i] = n_exact;
        if (overrun) atomicExch(&v.flags->overrun, 1);
        else if (!accepted) atomicExch(&v.flags->stalled, 1);
    }
}

cudaError_t launch_ess_theta(cudaStream_t st, const ThetaView& v) {
    if (v.n_local <= 0) return cudaSuccess;
    // few respondents on this rank (an item shard): four times the threads per respondent keep the SMs' memory pipes busy
    const bool wide = g_theta_threads ? g_theta_threads == 1024 : v.n_local <= 600;
    ThetaView w = v;
    const size_t row_bytes = ((size_t)v.m_total + 15) / 16 * 16;
    w.y_in_smem = row_bytes <= (wide ? (size_t)96 << 10 : (size_t)24 << 10) && row_bytes <= (size_t)v.ldm;  // keeps >= 2 wide / 8 narrow CTAs per SM
    const size_t smem = w.y_in_smem ? row_bytes : 0;
    if (wide && smem > ((size_t)32 << 10)) {  // static (~9 KB: virtual lanes, partials) + dynamic above the 48 KB default: the opt-in is per device (one
                                              // process may drive several), so it is set where it is needed; it costs microseconds
        cudaError_t ea = cudaFuncSetAttribute(k_ess_theta<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 << 10);
        if (ea != cudaSuccess) return ea;
    }
    if (wide) k_ess_theta<1024><<<v.n_local, 1024, smem, st>>>(w);
    else k_ess_theta<256><<<v.n_local, 256, smem, st>>>(w);
    ++g_kernel_launches;
    return cudaGetLastError();
}


// One CTA per respondent; the H sessions' log-likelihood sums are formed one after the other (each a block reduction in the
// fixed order of k_ess_theta) and added in session order, as the reference's `log_y += compute_ll_sparse(...)` loop does.
__global__ void __launch_bounds__(THETA_THREADS) k_ess_theta_panel(ThetaPanelView v) {
    __shared__ double s_part[THETA_THREADS / 32];
    const int i = blockIdx.x;
    if (i >= v.n) return;
    const int tid = threadIdx.x;
    const int C = v.C, m = v.m, H = v.H;

    auto loglik_h = [&](int h, double th) -> double {  // compute_ll_sparse of session h, src/draw-theta.cpp:4-24
        const ThetaSession& S = v.ses[h];
        const int8_t* __restrict__ yrow = S.y_rm + (size_t)i * v.ldm;
        const double* __restrict__ table = S.table;
        const int idx = grid_index(th);
        const double ts = grid_value(idx);
        auto term = [&](int J) -> double {
            const int y = yrow[J];
            if (y == 0) return 0.0;
            if (table) return table[((size_t)idx * m + J) * C + (y - 1)];
            const double fs = S.fstar[(size_t)idx * v.ld_chunk + J];
            const double mu = mu_of(ts, S.beta[3 * (size_t)J], S.beta[3 * (size_t)J + 1], S.beta[3 * (size_t)J + 2]);
            const double g = __dadd_rn(fs, mu);  // f* already holds mu*; the reference adds it again (:17)
            return cell_ll(g, S.thr[(size_t)J * (C + 1) + y - 1], S.thr[(size_t)J * (C + 1) + y], y == 1, y == C);
        };
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int J = tid;
        for (; J + 3 * THETA_THREADS < m; J += 4 * THETA_THREADS) {
            a0 += term(J);
            a1 += term(J + THETA_THREADS);
            a2 += term(J + 2 * THETA_THREADS);
            a3 += term(J + 3 * THETA_THREADS);
        }
        for (; J < m; J += THETA_THREADS) a0 += term(J);
        return block_sum<THETA_THREADS / 32>((a0 + a1) + (a2 + a3), s_part);
    };

    // nu = L_time z, dense column sweep (rmvnorm_threadsafe: L * z, src/mvnormal.h:17-24)
    double nu[GPIRT_MAX_HORIZON], th0[GPIRT_MAX_HORIZON], raw[GPIRT_MAX_HORIZON];
    for (int h = 0; h < H; ++h) { nu[h] = 0.0; th0[h] = v.ses[h].theta[i]; raw[h] = th0[h]; }
    for (int k = 0; k < H; ++k) {
        const double zk = v.z ? v.z[(size_t)i * v.z_stride + k] : philox_normal(v.key, ST_THETA_Z, i, k);
        for (int h = 0; h < H; ++h) nu[h] = __dadd_rn(nu[h], __dmul_rn(v.L_time[h + (size_t)k * H], zk));
    }
    int overrun = 0;
    uint32_t k = 0;
    const double u = v.us.get(v.key, ST_THETA_U, i, k++, &overrun);
    double log_y = log(u);
    for (int h = 0; h < H; ++h) log_y = __dadd_rn(log_y, loglik_h(h, th0[h]));  // :43-47
    double eps_min = 0.0, eps_max = TWO_PI_R;
    double eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ST_THETA_U, i, k++, &overrun)));
    eps_min = __dsub_rn(eps, TWO_PI_R);
    int evals = 1;
    double margin = CUDART_INF;
    bool accepted = false;
    for (int it = 0; it < ESS_MAX_PROPOSALS; ++it) {
        const double ce = cos(eps), se = sin(eps);
        double llp = 0.0;
        for (int h = 0; h < H; ++h) {
            raw[h] = __dadd_rn(__dmul_rn(th0[h], ce), __dmul_rn(nu[h], se));
            raw[h] = fmax(-5.0, fmin(5.0, raw[h]));  // :58-62
        }
        for (int h = 0; h < H; ++h) llp = __dadd_rn(llp, loglik_h(h, raw[h]));
        ++evals;
        margin = fmin(margin, fabs(llp - log_y));
        if (llp > log_y) { accepted = true; break; }
        if (eps < 0.0) eps_min = eps; else eps_max = eps;
        eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ST_THETA_U, i, k++, &overrun)));
        if (overrun) break;
    }
    if (tid == 0) {
        for (int h = 0; h < H; ++h) v.ses[h].theta_new[i] = grid_value(grid_index(raw[h]));  // :229-233
        v.evals[i] = evals;
        v.margin[i] = margin;
        if (overrun) atomicExch(&v.flags->overrun, 1);
        else if (!accepted) atomicExch(&v.flags->stalled, 1);
    }
}

cudaError_t launch_ess_theta_panel(cudaStream_t st, const ThetaPanelView& v) {
    if (v.n <= 0) return cudaSuccess;
    k_ess_theta_panel<<<v.n, THETA_THREADS, 0, st>>>(v);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
__global__ void k_gather_f(const double* __restrict__ theta, int n, const double* __restrict__ fstar, int ld_chunk,
                           int m_local, double* __restrict__ f, int ldn) {
    __shared__ double t[32][33];
    const int i0 = blockIdx.x * 32, j0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r, j = j0 + tx;
        if (i < n && j < m_local) t[r][tx] = fstar[(size_t)grid_index(theta[i]) * ld_chunk + j];
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int j = j0 + r, i = i0 + tx;
        if (i < n && j < m_local) f[(size_t)j * ldn + i] = t[tx][r];
    }
}

cudaError_t launch_gather_f(cudaStream_t st, const double* theta, int n, const double* fstar_chunk, int ld_chunk, int m_local,
                            double* f, int ldn) {
    dim3 grid((n + 31) / 32, (m_local + 31) / 32), blk(32, 8);
    k_gather_f<<<grid, blk, 0, st>>>(theta, n, fstar_chunk, ld_chunk, m_local, f, ldn);
    ++g_kernel_launches;
    return cudaGetLastError();
}

__global__ void k_add_mustar(double* __restrict__ fstar, int ld_chunk, int m_local, int item0, const double* __restrict__ beta, int p_beta) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int idx = blockIdx.y;
    if (j >= m_local) return;
    const size_t J = (size_t)item0 + j;
    const double ts = grid_value(idx);
    const double mu = mu_of(ts, beta[3 * J], beta[3 * J + 1], p_beta == 3 ? beta[3 * J + 2] : 0.0);
    fstar[(size_t)idx * ld_chunk + j] += mu;
}

cudaError_t launch_add_mustar(cudaStream_t st, double* fstar_chunk, int ld_chunk, int m_local, int item0, const double* beta, int p_beta) {
    dim3 grid((m_local + 127) / 128, NGRID);
    k_add_mustar<<<grid, 128, 0, st>>>(fstar_chunk, ld_chunk, m_local, item0, beta, p_beta);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// f* += mu* on the local chunk and, in the same pass, the finished values go to chunk `rank` of every peer's table (16-byte stores over
// NVLink; ld_chunk is a multiple of 16 doubles and the chunk is 256-byte aligned on every rank). One thread per pair of items of one grid row.
__global__ void __launch_bounds__(128) k_add_mustar_push(double* __restrict__ fstar, int ld_chunk, int m_local, int item0, const double* __restrict__ beta,
                                                         int p_beta, PeerPush peers) {
    const int j = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int idx = blockIdx.y;
    if (j >= m_local) return;
    const double ts = grid_value(idx);
    const size_t off = (size_t)idx * ld_chunk + j;
    double2 v = *reinterpret_cast<const double2*>(fstar + off);
    const size_t J = (size_t)item0 + j;
    v.x += mu_of(ts, beta[3 * J], beta[3 * J + 1], p_beta == 3 ? beta[3 * J + 2] : 0.0);
    if (j + 1 < m_local) v.y += mu_of(ts, beta[3 * J + 3], beta[3 * J + 4], p_beta == 3 ? beta[3 * J + 5] : 0.0);
    *reinterpret_cast<double2*>(fstar + off) = v;
#pragma unroll 1
    for (int p = 0; p < peers.count; ++p) *reinterpret_cast<double2*>(peers.chunk[p] + off) = v;
}

cudaError_t launch_add_mustar_push(cudaStream_t st, double* fstar_chunk, int ld_chunk, int m_local, int item0, const double* beta, int p_beta,
                                   const PeerPush& peers) {
    dim3 grid(((m_local + 1) / 2 + 127) / 128, NGRID);
    k_add_mustar_push<<<grid, 128, 0, st>>>(fstar_chunk, ld_chunk, m_local, item0, beta, p_beta, peers);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// Runs after k_add_mustar_push on the same stream: the pushed values are ordered before the flag by the kernel boundary and the system fence.
__global__ void k_signal_peers(PeerPush peers, unsigned long long epoch) {
    __threadfence_system();
    const int p = threadIdx.x;
    if (p < peers.count) {
        volatile unsigned long long* f = peers.flag[p];
        *f = epoch;
    }
    __threadfence_system();
}
cudaError_t launch_signal_peers(cudaStream_t st, const PeerPush& peers, unsigned long long epoch) {
    k_signal_peers<<<1, 32, 0, st>>>(peers, epoch);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// Holds the stream until every peer's flag word has reached `epoch` (their chunks of the f* table are then complete in this GPU's memory).
// Bounded: after ~4 s without progress it reports peer_timeout instead of hanging the device.
__global__ void k_wait_peers(const unsigned long long* flags, int world, int rank, unsigned long long epoch, DevFlags* dev_flags, int only) {
    const int p = threadIdx.x;
    if (p < world && p != rank && (only < 0 || p == only)) {
        const volatile unsigned long long* f = flags + p;
        unsigned long long t0 = 0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (*f < epoch) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 4000000000ULL) {
                atomicExch(&dev_flags->peer_timeout, p + 1);
                break;
            }
            __nanosleep(200);
        }
    }
    __threadfence_system();
}
cudaError_t launch_wait_peers(cudaStream_t st, const unsigned long long* flags, int world, int rank, unsigned long long epoch, DevFlags* dev_flags, int only) {
    k_wait_peers<<<1, 32, 0, st>>>(flags, world, rank, epoch, dev_flags, only);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// one thread per pair of rows of one column; 16-byte loads and stores where the pair is whole
__global__ void __launch_bounds__(128) k_push_block(const double* __restrict__ src, int ld, int rows, int cols, PeerBlockPush pb) {
    const int r = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int c = blockIdx.y;
    if (r >= rows || c >= cols) return;
    const size_t off = (size_t)c * ld + r;
    if (r + 1 < rows) {
        const double2 v = *reinterpret_cast<const double2*>(src + off);
#pragma unroll 1
        for (int p = 0; p < pb.count; ++p) *reinterpret_cast<double2*>(pb.dst[p] + off) = v;
    } else {
        const double v = src[off];
#pragma unroll 1
        for (int p = 0; p < pb.count; ++p) pb.dst[p][off] = v;
    }
}
cudaError_t launch_push_block(cudaStream_t st, const double* src, int ld, int rows, int cols, const PeerBlockPush& pb) {
    if (rows <= 0 || cols <= 0 || pb.count <= 0) return cudaSuccess;
    dim3 grid(((rows + 1) / 2 + 127) / 128, cols);
    k_push_block<<<grid, 128, 0, st>>>(src, ld, rows, cols, pb);
    ++g_kernel_launches;
    return cudaGetLastError();
}

__global__ void k_ordered_sum(const double* __restrict__ x, int n, double* out) {
    // fixed-shape reduction: 256 strided partial sums, combined in index order by thread 0
    __shared__ double s[256];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) acc += x[i];
    s[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < 256; ++k) t += s[k];
        *out = t;
    }
}
cudaError_t launch_ordered_sum(cudaStream_t st, const double* x, int n, double* out) {
    k_ordered_sum<<<1, 256, 0, st>>>(x, n, out);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// out[c + r*ld_out] = in[r + c*ld_in]
__global__ void k_transpose(const double* __restrict__ in, int ld_in, int rows, int cols, double* __restrict__ out, int ld_out) {
    __shared__ double t[32][33];
    const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int k = ty; k < 32; k += 8) {
        const int r = r0 + tx, c = c0 + k;
        if (r < rows && c < cols) t[k][tx] = in[(size_t)c * ld_in + r];
    }
    __syncthreads();
    for (int k = ty; k < 32; k += 8) {
        const int c = c0 + tx, r = r0 + k;
        if (r < rows && c < cols) out[(size_t)r * ld_out + c] = t[tx][k];
    }
}
cudaError_t launch_transpose(cudaStream_t st, const double* in, int ld_in, int rows, int cols, double* out, int ld_out) {
    dim3 grid((rows + 31) / 32, (cols + 31) / 32), blk(32, 8);
    k_transpose<<<grid, blk, 0, st>>>(in, ld_in, rows, cols, out, ld_out);
    ++g_kernel_launches;
    return cudaGetLastError();
}

__global__ void k_clamp_theta(double* theta, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) theta[i] = fmax(-5.0, fmin(5.0, theta[i]));  // theta.clamp(-5,5), src/gpirtMCMC.cpp:133
}
cudaError_t launch_clamp_theta(cudaStream_t st, double* theta, int n) {
    k_clamp_theta<<<(n + 255) / 256, 256, 0, st>>>(theta, n);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// y (float64, NaN missing, codes 1..C) -> int8 (0 missing). Non-finite == missing (arma::find_finite).
__global__ void k_y_to_i8(const double* __restrict__ y, long long count, int C, int8_t* __restrict__ out, DevFlags* flags) {
    long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= count) return;
    const double v = y[e];
    int8_t code = 0;
    if (isfinite(v)) {
        const int c = (int)v;  // int(y[i]), src/log-likelihood.cpp:57
        if (c < 1 || c > C) { atomicExch(&flags->bad_code, 1); code = 0; }
        else code = (int8_t)c;
    }
    out[e] = code;
}
cudaError_t launch_y_to_i8(cudaStream_t st, const double* y, long long count, int C, int8_t* out, DevFlags* flags) {
    k_y_to_i8<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(y, count, C, out, flags);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// From the full n x m_total int8 matrix (column-major) build the two resident layouts:
//   y_im[j*ldy + i]   item-major columns of the local items (padding zero = missing)
//   y_rm[il*ldm + J]  respondent-major rows of the local respondents
__global__ void k_y_layouts(const int8_t* __restrict__ y, int n, int m_total, int item0, int m_local, int8_t* __restrict__ y_im, int ldy,
                            int resp0, int n_local, int8_t* __restrict__ y_rm, long long ldm) {
    __shared__ int8_t t[32][33];
    const int i0 = blockIdx.x * 32, J0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int k = ty; k < 32; k += 8) {
        const int i = i0 + tx, J = J0 + k;
        int8_t val = (i < n && J < m_total) ? y[(size_t)J * n + i] : 0;
        t[k][tx] = val;
        if (i < n && J >= item0 && J < item0 + m_local) y_im[(size_t)(J - item0) * ldy + i] = val;
    }
    __syncthreads();
    for (int k = ty; k < 32; k += 8) {
        const int J = J0 + tx, i = i0 + k;
        if (J < m_total && i >= resp0 && i < resp0 + n_local) y_rm[(size_t)(i - resp0) * ldm + J] = t[tx][k];
    }
}
cudaError_t launch_y_layouts(cudaStream_t st, const int8_t* y_full, int n, int m_total, int item0, int m_local, int8_t* y_im, int ldy,
                             int resp0, int n_local, int8_t* y_rm, long long ldm) {
    dim3 grid((n + 31) / 32, (m_total + 31) / 32), blk(32, 8);
    k_y_layouts<<<grid, blk, 0, st>>>(y_full, n, m_total, item0, m_local, y_im, ldy, resp0, n_local, y_rm, ldm);
    ++g_kernel_launches;
    return cudaGetLastError();
}

}  // namespace gpirt

namespace gpirt {

// ------------------------------------------------------------------------------------------------
// Item-sharded theta step (samplers.cuh, ThetaItemsView).
// One ESS unit's proposal sequence under rejections: the arithmetic of k_ess_theta, operation for operation (src/draw-theta.cpp:38-62).
struct ThetaWalk {
    double th0, nu, log_u, eps, eps_min, eps_max;
    uint32_t k;      // uniforms consumed
    int overrun;
    __device__ __forceinline__ void init(const ThetaItemsView& v, int i) {
        double Li;
        if (v.regime == 0) {
            const double sd = v.theta_prior_sds[2 * (size_t)i];
            Li = sqrt(__dadd_rn(1.0, __dmul_rn(sd, sd)));
        } else {
            Li = sqrt(__dadd_rn(__dmul_rn(v.theta_os, v.theta_os), 1e-6));
        }
        const double z = v.z ? v.z[(size_t)i * v.z_stride] : philox_normal(v.key, ST_THETA_Z, i, 0);
        nu = __dmul_rn(Li, z);
        th0 = v.theta[i];
        overrun = 0;
        k = 0;
        log_u = log(v.us.get(v.key, ST_THETA_U, i, k++, &overrun));
        eps_min = 0.0; eps_max = TWO_PI_R;
        eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ST_THETA_U, i, k++, &overrun)));
        eps_min = __dsub_rn(eps, TWO_PI_R);
    }
    __device__ __forceinline__ double proposal() const {
        const double raw = __dadd_rn(__dmul_rn(th0, cos(eps)), __dmul_rn(nu, sin(eps)));
        return fmax(-5.0, fmin(5.0, raw));
    }
    __device__ __forceinline__ void reject(const ThetaItemsView& v, int i) {
        if (eps < 0.0) eps_min = eps; else eps_max = eps;
        eps = __dadd_rn(eps_min, __dmul_rn(__dsub_rn(eps_max, eps_min), v.us.get(v.key, ST_THETA_U, i, k++, &overrun)));
    }
};

// One CTA per respondent: partial log-likelihoods over the local items of the current state and of the next THETA_BATCH proposals. The item loop is
// the outer one, so a thread keeps THETA_BATCH + 1 independent table gathers in flight per item instead of one evaluation after the other.
__global__ void __launch_bounds__(256) k_theta_partials(ThetaItemsView v) {
    constexpr int S = THETA_BATCH + 1, T = 256;
    __shared__ int s_idx[S];
    __shared__ double s_red[S][T / 32];
    extern __shared__ __align__(16) int8_t s_yl[];
    const int i = blockIdx.x, tid = threadIdx.x;
    const int k0 = v.k_next[i];
    if (k0 < 0) return;  // finished in an earlier round
    if (tid == 0) {
        ThetaWalk w;
        w.init(v, i);
        for (int q = 0; q < k0; ++q) w.reject(v, i);
        s_idx[0] = grid_index(w.th0);
        for (int b = 1; b < S; ++b) {
            s_idx[b] = grid_index(w.proposal());
            w.reject(v, i);  // a supplied stream may run dry here for proposals no decision will ever reach: only k_theta_decide reports an overrun
        }
    }
    const int m = v.m_local, C = v.C;
    const int8_t* yrow = v.y_rl + (size_t)i * v.ldml;
    const bool y_smem = (size_t)((m + 15) / 16 * 16) <= (size_t)24 << 10;
    if (y_smem) {
        const int4* src = reinterpret_cast<const int4*>(yrow);
        int4* dst = reinterpret_cast<int4*>(s_yl);
        for (int q = tid; q < (m + 15) / 16; q += T) dst[q] = src[q];
        yrow = s_yl;
    }
    __syncthreads();
    int idx[S];
#pragma unroll
    for (int b = 0; b < S; ++b) idx[b] = s_idx[b];
    const int b_first = k0 == 0 ? 0 : 1;  // the current state is scored in the first round only
    double acc[S];
#pragma unroll
    for (int b = 0; b < S; ++b) acc[b] = 0.0;
    const size_t row = (size_t)m * C;
    for (int J = tid; J < m; J += T) {
        const int y = yrow[J];
        if (y == 0) continue;
        const double* base = v.table + (size_t)J * C + (y - 1);
#pragma unroll
        for (int b = 0; b < S; ++b)
            if (b >= b_first) acc[b] += base[(size_t)idx[b] * row];
    }
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int b = 0; b < S; ++b) {
        const double w = warp_sum(acc[b]);
        if (lane == 0) s_red[b][warp] = w;
    }
    __syncthreads();
    if (tid < S) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < T / 32; ++w) t += s_red[tid][w];
        v.part_mine[(size_t)i * S + tid] = t;
    }
}

cudaError_t launch_theta_partials(cudaStream_t st, const ThetaItemsView& v) {
    if (v.n <= 0) return cudaSuccess;
    const size_t row_bytes = ((size_t)v.m_local + 15) / 16 * 16;
    const size_t smem = row_bytes <= ((size_t)24 << 10) ? row_bytes : 0;
    k_theta_partials<<<v.n, 256, smem, st>>>(v);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// One thread per respondent, the same on every rank: partials added in rank order, then the reference's accept test proposal by proposal.
__global__ void __launch_bounds__(128) k_theta_decide(ThetaItemsView v) {
    constexpr int S = THETA_BATCH + 1;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= v.n) return;
    const int k0 = v.k_next[i];
    if (k0 < 0) return;
    auto total = [&](int slot) {
        double t = 0.0;
        for (int r = 0; r < v.world; ++r) t += v.part_all[((size_t)r * v.n + i) * S + slot];
        return t;
    };
    ThetaWalk w;
    w.init(v, i);
    for (int q = 0; q < k0; ++q) w.reject(v, i);
    if (k0 == 0) { v.ll0[i] = total(0); v.margin[i] = CUDART_INF; }
    const double log_y = __dadd_rn(w.log_u, v.ll0[i]);
    double margin = v.margin[i];
    double raw = w.th0;
    int done = 0, over = w.overrun;
    int kk = k0;
    for (int b = 1; b < S && !over; ++b) {
        raw = w.proposal();
        const double llp = total(b);
        ++kk;
        margin = fmin(margin, fabs(llp - log_y));
        if (llp > log_y) { done = 1; break; }
        if (kk >= ESS_MAX_PROPOSALS) { done = 2; break; }
        w.reject(v, i);
        over = w.overrun;
    }
    v.margin[i] = margin;
    if (over) { atomicExch(&v.flags->overrun, 1); done = 3; }
    if (done) {
        v.theta_new[i] = grid_value(grid_index(raw));  // snapped to the grid, src/draw-theta.cpp:165-169 (on a stall: the last proposal, as k_ess_theta)
        v.evals[i] = 1 + kk;
        v.k_next[i] = -1;
        if (done == 2) atomicExch(&v.flags->stalled, 1);
    } else {
        v.k_next[i] = kk;
        atomicAdd(v.n_active, 1);
    }
}

cudaError_t launch_theta_decide(cudaStream_t st, const ThetaItemsView& v) {
    if (v.n <= 0) return cudaSuccess;
    k_theta_decide<<<(v.n + 127) / 128, 128, 0, st>>>(v);
    ++g_kernel_launches;
    return cudaGetLastError();
}

// (beta 3 x m_chunk | thresholds (C+1) x m_chunk | item ll m_chunk) of every rank's chunk <-> packed[world][m_chunk * (C1 + 4)]; pack touches this
// rank's chunk only, unpack every other rank's
__global__ void k_pack_items(double* __restrict__ beta, double* __restrict__ thr, double* __restrict__ item_ll, int C1, int m_chunk, int rank, int world,
                             double* __restrict__ packed, int unpack) {
    const int per = m_chunk * (C1 + 4);
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)per * world) return;
    const int r = (int)(e / per), q = (int)(e - (long long)r * per);
    if (unpack ? r == rank : r != rank) return;
    double* p;
    if (q < 3 * m_chunk) p = beta + (size_t)r * 3 * m_chunk + q;
    else if (q < (3 + C1) * m_chunk) p = thr + (size_t)r * C1 * m_chunk + (q - 3 * m_chunk);
    else p = item_ll + (size_t)r * m_chunk + (q - (3 + C1) * m_chunk);
    if (unpack) *p = packed[e]; else packed[e] = *p;
}
cudaError_t launch_pack_items(cudaStream_t st, const double* beta, const double* thr, const double* item_ll, int C1, int m_chunk, int rank, int world,
                              double* packed, int unpack) {
    const long long total = (long long)m_chunk * (C1 + 4) * world;
    k_pack_items<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(const_cast<double*>(beta), const_cast<double*>(thr), const_cast<double*>(item_ll), C1, m_chunk,
                                                                rank, world, packed, unpack);
    ++g_kernel_launches;
    return cudaGetLastError();
}

}  // namespace gpirt
