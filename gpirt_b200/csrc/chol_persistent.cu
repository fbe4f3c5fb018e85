// Lower Cholesky of the covariance matrices of the sweep (arma::chol(S, "lower"): src/cholesky-cache.cpp:23 for K(theta, theta) + 1e-6 I,
// src/draw-fstar.cpp:36 for K_post) as ONE persistent kernel.
//
// Why: the factorisation is a chain of n / 64 panel steps, each potf2(64 x 64) -> panel solve -> trailing update. As separate launches
// (linalg.cu, chol_lower_blocked) a step costs ~37 us at n = 2000 — the redundant potf2 in every CTA, the solve, a latency-bound K = 64 GEMM
// launch — and the whole chain sits on the critical path of the sweep. Here the matrix is cut into 64 x 64 tiles and the steps overlap:
//   * one CTA (the first to arrive) walks the critical path: potf2 of the diagonal tile (k, k), then IN THE SAME CTA the solve of tile
//     (k+1, k) and the update of tile (k+1, k+1), which it keeps in shared memory as the next step's diagonal tile — no global round trip,
//     no launch, no grid-wide barrier between two panel steps;
//   * every other CTA takes tile tasks from one ordered list through an atomic ticket: S(i, k) = solve of panel tile (i, k) against L(k, k),
//     U(i, j, k) = tile (i, j) -= L(i, k) L(j, k)^T on the fp64 tensor pipe (DMMA.8x8x4). A task waits for its inputs on per-tile version
//     counters in global memory (tile (i, j) carries the number of updates applied; it is final at version j + 1). Every dependency of a
//     task is produced by the critical CTA or by a task EARLIER in the list, so whatever number of CTAs is resident the lowest outstanding
//     ticket can always run: no deadlock, no cooperative launch. The updates of one tile are applied in step order, so the factor does not
//     depend on scheduling (bit-reproducible).
//   * tiles are staged global -> shared with TMA bulk copies (cp.async.bulk + mbarrier complete_tx, one 512-byte column per copy), issued
//     ahead of use so the next tile lands while the current one is being factored.
// Spins are bounded (2 s on %globaltimer): a stall raises `abort`, every CTA leaves, the host reports an error instead of hanging the GPU.
#include <cstdint>
#include <cstdlib>
#include <vector>

#include "linalg.cuh"

namespace gpirt {

namespace {

constexpr int TB_ = 64;          // tile edge
constexpr int TP = 72;           // shared-memory pitch of a tile column (doubles): == 8 mod 16, DMMA fragment loads take the minimal two wavefronts
constexpr int TILE_ELEMS = TB_ * TP;
constexpr int CP_THREADS = 256;
constexpr unsigned long long SPIN_LIMIT_NS = 2000000000ULL;

// task word: type (2 bits) | i (10) | j (10) | k (10)
enum : uint32_t { TASK_S = 1u, TASK_U = 2u };
__host__ __device__ inline uint32_t pack_task(uint32_t type, int i, int j, int k) { return (type << 30) | ((uint32_t)i << 20) | ((uint32_t)j << 10) | (uint32_t)k; }

struct Shared {
    double A[TILE_ELEMS];
    double B[TILE_ELEMS];
    double C[TILE_ELEMS];
    double invd[TB_];
    unsigned long long mbar[3];
    uint32_t task;
    int role;
    int abort;
};

// ---- PTX wrappers: mbarrier + TMA bulk copy (async proxy) + acquire / release flags ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned long long globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

struct Ctx {
    double* A; int ld; int n; int T;
    int* state;          // [0] ticket, [1] arrivals, [2] abort, [3] unused, [4 + i*T + j] version of tile (i, j)
    int* info; int info_base;
    Shared* sh;
    uint32_t par[3];     // phase parity of the three mbarriers (thread-uniform bookkeeping)
    long long* trace;    // optional clock64 stamps (gpirt_debug_chol_trace): [k * 16 + s] for the critical CTA, [T * 16 + t * 8 + s] for worker 1
    __device__ int* ver(int i, int j) const { return state + 4 + i * T + j; }
};

// thread 0 spins until *p >= want (or the abort flag / time limit); every thread returns the same verdict
__device__ __forceinline__ bool wait_ge(Ctx& c, const int* p, int want) {
    if (threadIdx.x == 0) {
        if (ld_acquire(p) < want) {
            const unsigned long long t0 = globaltimer();
            while (ld_acquire(p) < want) {
                __nanosleep(32);
                if (ld_acquire(c.state + 2) != 0) { c.sh->abort = 1; break; }
                if (globaltimer() - t0 > SPIN_LIMIT_NS) { atomicExch(c.state + 2, 1); c.sh->abort = 1; break; }
            }
        }
        fence_proxy_async();  // the tiles behind the flag are read next through the async proxy (TMA)
    }
    __syncthreads();
    return c.sh->abort == 0;
}
#define TRACE(slot) do { if (c.trace && threadIdx.x == 0) c.trace[(slot)] = clock64(); } while (0)
__device__ __forceinline__ bool ready(const int* p, int want) { return ld_acquire(p) >= want; }

// all threads have written the tile to global memory: publish its new version
__device__ __forceinline__ void publish(Ctx& c, int* p, int v) {
    __syncthreads();
    if (threadIdx.x == 0) st_release(p, v);  // release at gpu scope is cumulative over the stores the barrier ordered before this thread
}

// TMA: columns [0, cols) of the tile at g (leading dimension ld), rows_pad rows (even, within the zero-padded leading dimension) -> sm, column pitch TP.
// Called by every thread of the CTA: each one fences its own generic-proxy accesses of the buffer against the async proxy, thread 0 arms the
// buffer's mbarrier with the byte count, the barrier ends all use of `sm`, then every warp issues 8 of the 64 column copies (one per lane): the
// bulk-copy instruction runs on the warp's uniform datapath, and 64 of them from one warp measured ~4.5k cycles of issue per tile (clock64 trace,
// profiles/r02_chol_persistent.txt) on the critical path.
__device__ __forceinline__ void tile_load(Ctx& c, int which, double* sm, const double* g, int rows_pad, int cols) {
    unsigned long long* bar = &c.sh->mbar[which];
    fence_proxy_async();
    if (threadIdx.x == 0) mbar_expect_tx(bar, (uint32_t)(rows_pad * 8 * cols));
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane < 8) {
        const int col = warp + 8 * lane;
        if (col < cols) bulk_g2s(sm + col * TP, g + (size_t)col * c.ld, (uint32_t)(rows_pad * 8), bar);
    }
}
__device__ __forceinline__ bool tile_wait(Ctx& c, int which) {
    unsigned long long* bar = &c.sh->mbar[which];
    const uint32_t parity = c.par[which];
    c.par[which] ^= 1u;
    if (!mbar_try_wait(bar, parity)) {
        const unsigned long long t0 = globaltimer();
        while (!mbar_try_wait(bar, parity)) {
            if (globaltimer() - t0 > SPIN_LIMIT_NS) { atomicExch(c.state + 2, 1); c.sh->abort = 1; break; }
        }
    }
    return true;
}
__device__ __forceinline__ int rows_of(const Ctx& c, int i) { return min(TB_, c.n - i * TB_); }
__device__ __forceinline__ int pad2(int r) { return (r + 1) & ~1; }
__device__ __forceinline__ const double* tile_ptr(const Ctx& c, int i, int j) { return c.A + (size_t)j * TB_ * c.ld + (size_t)i * TB_; }

__device__ __forceinline__ void sqrt_rsqrt2(double d, double& s, double& r) {  // sqrt(d) and 1/sqrt(d) without a division
    r = rsqrt(d);
    s = d * r;
    s = fma(fma(-s, s, d), 0.5 * r, s);
}
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// ---- potf2 of the kb x kb diagonal tile held in sm (lower part; element (r, c) at sm[c * TP + r]) ----
// Eight micro-steps of eight columns: (a) one thread factors the 8 x 8 diagonal block in registers (the pivot chain: rsqrt + Newton per column, no
// barrier inside), (b) the rows below solve their 8 columns against it (one thread per row), (c) the rest of the tile takes the rank-8 update on
// the DMMA pipe (one 8 x 8 block = two mma.m8n8k4). Three barriers per eight columns.
__device__ void potf2_tile(Ctx& c, double* sm, int kb) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, grp = lane >> 2, tig = lane & 3;
    // rows / columns beyond kb (ragged last tile): identity, so the padded part factors to itself
    for (int e = tid; e < TB_ * TB_; e += CP_THREADS) {
        const int r = e & 63, col = e >> 6;
        if (r >= kb || col >= kb) sm[col * TP + r] = (r == col) ? 1.0 : 0.0;
    }
    __syncthreads();
    for (int s = 0; s < 8; ++s) {
        const int c0 = 8 * s;
        if (tid == 0) {
            double a[8][8];
#pragma unroll
            for (int col = 0; col < 8; ++col)
#pragma unroll
                for (int r = col; r < 8; ++r) a[r][col] = sm[(c0 + col) * TP + c0 + r];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                double d = a[j][j];
                if (!(d > 0.0)) {  // arma::chol throws here: record the failing leading minor, keep the arithmetic finite
                    atomicCAS(c.info, 0, c.info_base + c0 + j + 1);
                    d = 1.0;
                }
                double sq, rs;
                sqrt_rsqrt2(d, sq, rs);
                a[j][j] = sq;
                c.sh->invd[c0 + j] = rs;
#pragma unroll
                for (int i = j + 1; i < 8; ++i) a[i][j] *= rs;
#pragma unroll
                for (int jj = j + 1; jj < 8; ++jj)
#pragma unroll
                    for (int i = jj; i < 8; ++i) a[i][jj] -= a[i][j] * a[jj][j];
            }
#pragma unroll
            for (int col = 0; col < 8; ++col)
#pragma unroll
                for (int r = col; r < 8; ++r) sm[(c0 + col) * TP + c0 + r] = a[r][col];
        }
        __syncthreads();
        if (s == 7) break;
        if (tid >= c0 + 8 && tid < TB_) {  // (b) row tid: x D^T = a over the 8 columns
            const int r = tid;
            double x[8];
#pragma unroll
            for (int col = 0; col < 8; ++col) x[col] = sm[(c0 + col) * TP + r];
#pragma unroll
            for (int col = 0; col < 8; ++col) {
#pragma unroll
                for (int k = 0; k < col; ++k) x[col] -= x[k] * sm[(c0 + k) * TP + c0 + col];
                x[col] *= c.sh->invd[c0 + col];
            }
#pragma unroll
            for (int col = 0; col < 8; ++col) sm[(c0 + col) * TP + r] = x[col];
        }
        __syncthreads();
        // (c) blocks (bi, bj), s < bj <= bi < 8: tile(bi, bj) -= X(bi) X(bj)^T
        int cnt = 0;
        for (int bi = s + 1; bi < 8; ++bi)
            for (int bj = s + 1; bj <= bi; ++bj, ++cnt) {
                if ((cnt & 7) != warp) continue;
                double* cp = sm + (8 * bj + 2 * tig) * TP + 8 * bi + grp;
                double v0 = cp[0], v1 = cp[TP];
#pragma unroll
                for (int kk = 0; kk < 8; kk += 4) {
                    const double af = -sm[(c0 + kk + tig) * TP + 8 * bi + grp];
                    const double bf = sm[(c0 + kk + tig) * TP + 8 * bj + grp];
                    dmma(v0, v1, af, bf);
                }
                cp[0] = v0;
                cp[TP] = v1;
            }
        __syncthreads();
    }
}

// ---- X L^T = A for the 64-row tile in sx against the factored diagonal tile in sl; invd = 1 / diag(L) ----
__device__ void solve_tile(Ctx& c, const double* sl, double* sx) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, grp = lane >> 2, tig = lane & 3;
    for (int s = 0; s < 8; ++s) {
        const int c0 = 8 * s;
        if (tid < TB_) {
            const int r = tid;
            double x[8];
#pragma unroll
            for (int col = 0; col < 8; ++col) x[col] = sx[(c0 + col) * TP + r];
#pragma unroll
            for (int col = 0; col < 8; ++col) {
#pragma unroll
                for (int k = 0; k < col; ++k) x[col] -= x[k] * sl[(c0 + k) * TP + c0 + col];
                x[col] *= c.sh->invd[c0 + col];
            }
#pragma unroll
            for (int col = 0; col < 8; ++col) sx[(c0 + col) * TP + r] = x[col];
        }
        __syncthreads();
        if (s == 7) break;
        // rows 8 warp .. 8 warp + 7: columns beyond the block take X(:, c0..c0+7) L(cols, c0..c0+7)^T
        double af[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) af[h] = -sx[(c0 + 4 * h + tig) * TP + 8 * warp + grp];
        for (int cb = s + 1; cb < 8; ++cb) {
            double* cp = sx + (8 * cb + 2 * tig) * TP + 8 * warp + grp;
            double v0 = cp[0], v1 = cp[TP];
#pragma unroll
            for (int h = 0; h < 2; ++h) dmma(v0, v1, af[h], sl[(c0 + 4 * h + tig) * TP + 8 * cb + grp]);
            cp[0] = v0;
            cp[TP] = v1;
        }
        __syncthreads();
    }
}

// acc(64 x 64) = sa sb^T over K = 64: 2 x 4 warps, each 32 x 16 of the tile
__device__ __forceinline__ void tile_product(const double* sa, const double* sb, double (&acc)[4][2][2]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, grp = lane >> 2, tig = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
#pragma unroll 4
    for (int kk = 0; kk < TB_; kk += 4) {
        double af[4], bf[2];
#pragma unroll
        for (int i = 0; i < 4; ++i) af[i] = sa[(kk + tig) * TP + 32 * wm + 8 * i + grp];
#pragma unroll
        for (int j = 0; j < 2; ++j) bf[j] = sb[(kk + tig) * TP + 16 * wn + 8 * j + grp];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
}

__device__ __forceinline__ void load_invd(Ctx& c, const double* sl) {  // 1 / L(r, r) without an fp64 division
    if (threadIdx.x < TB_) {
        const double v = sl[threadIdx.x * TP + threadIdx.x];
        c.sh->invd[threadIdx.x] = rsqrt(v * v);
    }
}

// store rows [0, rows) x columns [0, cols) of the shared tile to global; lower_zero: the diagonal tile, zeros above the diagonal
__device__ __forceinline__ void tile_store(Ctx& c, const double* sm, int i, int j, int rows, int cols, bool diag) {
    double* g = const_cast<double*>(tile_ptr(c, i, j));
    for (int e = threadIdx.x; e < TB_ * TB_; e += CP_THREADS) {
        const int r = e & 63, col = e >> 6;
        if (r < rows && col < cols) g[(size_t)col * c.ld + r] = (diag && col > r) ? 0.0 : sm[col * TP + r];
    }
}

// ---- the critical path: potf2(k, k), solve (k+1, k), update (k+1, k+1) -> next diagonal tile, all in this CTA ----
__device__ void run_critical(Ctx& c) {
    Shared* sh = c.sh;
    const int T = c.T;
    tile_load(c, 0, sh->A, tile_ptr(c, 0, 0), pad2(rows_of(c, 0)), rows_of(c, 0));
    bool diag_in_flight = true;
    for (int k = 0; k < T; ++k) {
        const int kb = rows_of(c, k);
        const bool has_next = k + 1 < T;
        const int nb = has_next ? rows_of(c, k + 1) : 0;
        TRACE(k * 16 + 0);
        // the panel tile (k+1, k) must carry the updates of steps < k; if it does already, its copy runs under the potf2
        bool sub_issued = false;
        if (has_next) {
            __syncthreads();
            if (threadIdx.x == 0) sh->task = ready(c.ver(k + 1, k), k) ? 1u : 0u;
            __syncthreads();
            if (sh->task) {
                tile_load(c, 1, sh->B, tile_ptr(c, k + 1, k), pad2(nb), TB_);
                sub_issued = true;
            }
        }
        if (diag_in_flight) { tile_wait(c, 0); diag_in_flight = false; }
        if (sh->abort) return;
        TRACE(k * 16 + 1);
        c.info_base = k * TB_;
        potf2_tile(c, sh->A, kb);
        TRACE(k * 16 + 2);
        tile_store(c, sh->A, k, k, kb, kb, true);
        publish(c, c.ver(k, k), k + 1);
        TRACE(k * 16 + 3);
        if (!has_next) break;
        if (!sub_issued) {
            if (!wait_ge(c, c.ver(k + 1, k), k)) return;
            tile_load(c, 1, sh->B, tile_ptr(c, k + 1, k), pad2(nb), TB_);
        }
        // the next diagonal tile (k+1, k+1) with the updates of steps < k: its copy runs under the solve when it is ready
        bool next_issued = false;
        __syncthreads();
        if (threadIdx.x == 0) sh->task = ready(c.ver(k + 1, k + 1), k) ? 1u : 0u;
        __syncthreads();
        if (sh->task) {
            tile_load(c, 2, sh->C, tile_ptr(c, k + 1, k + 1), pad2(nb), nb);
            next_issued = true;
        }
        TRACE(k * 16 + 4);
        tile_wait(c, 1);
        if (sh->abort) return;
        TRACE(k * 16 + 5);
        // invd of this tile is what potf2 left (1 / sqrt(pivot)): solve in place
        solve_tile(c, sh->A, sh->B);
        TRACE(k * 16 + 6);
        tile_store(c, sh->B, k + 1, k, nb, TB_, false);
        publish(c, c.ver(k + 1, k), k + 1);
        TRACE(k * 16 + 7);
        if (!next_issued) {
            if (!wait_ge(c, c.ver(k + 1, k + 1), k)) return;
            tile_load(c, 2, sh->C, tile_ptr(c, k + 1, k + 1), pad2(nb), nb);
        }
        TRACE(k * 16 + 8);
        tile_wait(c, 2);
        if (sh->abort) return;
        TRACE(k * 16 + 9);
        // next diagonal tile = C - X X^T, written over A (L(k, k) is no longer needed)
        {
            double acc[4][2][2];
            tile_product(sh->B, sh->B, acc);
            __syncthreads();
            const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, grp = lane >> 2, tig = lane & 3;
            const int wm = warp & 1, wn = warp >> 1;
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int r = 32 * wm + 8 * i + grp, col = 16 * wn + 8 * j + 2 * tig + e;
                        sh->A[col * TP + r] = sh->C[col * TP + r] - acc[i][j][e];
                    }
        }
        __syncthreads();
        TRACE(k * 16 + 10);
        // tile (k+1, k+1) now carries step k as well; nobody else reads it before its potf2 publishes version k + 2
    }
}

// ---- the other CTAs: tile tasks ----
// One task: S = solve of panel tile (i, k) against L(k, k); U = tile (i, j) -= L(i, k) L(j, k)^T. Returns false when the kernel is giving up.
__device__ bool do_task(Ctx& c, uint32_t w, int tb) {
    Shared* sh = c.sh;
    const int type = (int)(w >> 30), i = (int)((w >> 20) & 1023u), j = (int)((w >> 10) & 1023u), k = (int)(w & 1023u);
    const int ri = rows_of(c, i);
    TRACE(tb + 0);
    if (c.trace && threadIdx.x == 0) c.trace[tb + 7] = (long long)w;
    if (type == (int)TASK_S) {  // tile (i, k): X L(k, k)^T = A(i, k)
        if (!wait_ge(c, c.ver(k, k), k + 1)) return false;
        tile_load(c, 0, sh->A, tile_ptr(c, k, k), TB_, TB_);
        if (!wait_ge(c, c.ver(i, k), k)) return false;
        tile_load(c, 1, sh->B, tile_ptr(c, i, k), pad2(ri), TB_);
        TRACE(tb + 1);
        tile_wait(c, 0);
        __syncthreads();
        load_invd(c, sh->A);
        tile_wait(c, 1);
        __syncthreads();
        if (sh->abort) return false;
        TRACE(tb + 2);
        solve_tile(c, sh->A, sh->B);
        TRACE(tb + 3);
        tile_store(c, sh->B, i, k, ri, TB_, false);
        TRACE(tb + 4);
        publish(c, c.ver(i, k), k + 1);
        TRACE(tb + 5);
        return true;
    }
    const int rj = rows_of(c, j);
    if (!wait_ge(c, c.ver(i, k), k + 1)) return false;
    tile_load(c, 0, sh->A, tile_ptr(c, i, k), pad2(ri), TB_);
    const int lane = threadIdx.x & 31, grp = lane >> 2, tig = lane & 3;
    double* g = const_cast<double*>(tile_ptr(c, i, j));
    const double* sb = sh->A;
    if (j != i) {
        if (!wait_ge(c, c.ver(j, k), k + 1)) return false;
        tile_load(c, 1, sh->B, tile_ptr(c, j, k), pad2(rj), TB_);
        sb = sh->B;
    }
    if (!wait_ge(c, c.ver(i, j), k)) return false;
    TRACE(tb + 1);
    // the C tile goes through registers: loads issued before the product so their latency hides under it
    const int warp = threadIdx.x >> 5, wm = warp & 1, wn = warp >> 1;
    double cv[4][2][2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int r = 32 * wm + 8 * a + grp, col = 16 * wn + 8 * b + 2 * tig + e;
                cv[a][b][e] = (r < ri && col < rj) ? __ldcg(g + (size_t)col * c.ld + r) : 0.0;
            }
    tile_wait(c, 0);
    if (j != i) tile_wait(c, 1);
    __syncthreads();
    if (sh->abort) return false;
    TRACE(tb + 2);
    double acc[4][2][2];
    tile_product(sh->A, sb, acc);
    TRACE(tb + 3);
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int r = 32 * wm + 8 * a + grp, col = 16 * wn + 8 * b + 2 * tig + e;
                if (r < ri && col < rj) g[(size_t)col * c.ld + r] = cv[a][b][e] - acc[a][b][e];
            }
    TRACE(tb + 4);
    publish(c, c.ver(i, j), k + 1);
    TRACE(tb + 5);
    return true;
}

// Roles by order of arrival: 0 the critical path; 1 and 2 the two short chains the critical CTA's NEXT step waits for — S(k+2, k) then
// U(k+2, k+2, k), and U(k+2, k+1, k) — which must not queue behind bulk work (measured: in the common list they arrived ~5 us late at every
// step); >= 3 the bulk list through the ticket.
__device__ void run_worker(Ctx& c, const uint32_t* tasks, int n_tasks) {
    Shared* sh = c.sh;
    const int role = sh->role;
    if (role != 3) c.trace = nullptr;  // only the first bulk worker is traced
    if (role == 1 || role == 2) {
        for (int k = 0; k + 2 < c.T; ++k) {
            if (role == 1) {
                if (!do_task(c, pack_task(TASK_S, k + 2, k, k), 0)) return;
                if (!do_task(c, pack_task(TASK_U, k + 2, k + 2, k), 0)) return;
            } else {
                if (!do_task(c, pack_task(TASK_U, k + 2, k + 1, k), 0)) return;
            }
        }
        return;
    }
    for (int done = 0;; ++done) {
        if (done >= 256) c.trace = nullptr;
        __syncthreads();
        if (threadIdx.x == 0) {
            const int t = atomicAdd(c.state + 0, 1);
            sh->task = t < n_tasks ? tasks[t] : 0u;
        }
        __syncthreads();
        const uint32_t w = sh->task;
        if (w == 0u || sh->abort) return;
        if (!do_task(c, w, c.T * 16 + done * 8)) return;
    }
}

// The register budget is a template parameter: unbounded the kernel takes 174 registers (the 8 x 8 pivot block of one thread), and a CTA then leaves room
// for ONE 256-thread, 64-register ESS CTA on its SM for as long as the factorisation runs beside the item kernels; MINB = 2 compiles to 126 registers
// without spills (two ESS CTAs fit), MINB = 3 to 80 with 160 bytes of spills. GPIRT_CHOL_MINB selects (measured: profiles/r02_chol_persistent.txt).
template <int MINB>
__global__ void __launch_bounds__(CP_THREADS, MINB) k_chol_persistent(double* A, int ld, int n, int T, int* state, int* info, const uint32_t* tasks,
                                                                      int n_tasks, long long* trace) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Shared* sh = reinterpret_cast<Shared*>(smem_raw);
    Ctx c;
    c.A = A; c.ld = ld; c.n = n; c.T = T; c.state = state; c.info = info; c.info_base = 0; c.sh = sh; c.trace = trace;
    c.par[0] = c.par[1] = c.par[2] = 0u;
    if (threadIdx.x == 0) {
        for (int b = 0; b < 3; ++b) mbar_init(&sh->mbar[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        sh->abort = 0;
        sh->role = atomicAdd(state + 1, 1);  // the first CTA to run takes the critical path: it is, by construction, resident
    }
    __syncthreads();
    if (sh->role == 0) run_critical(c);
    else run_worker(c, tasks, n_tasks);
}

// tiles above the diagonal still hold K's upper half; and a stalled factorisation (abort flag) becomes info = -1 for the host
__global__ void k_zero_upper2(double* __restrict__ A, int ld, int n, const int* state, int* info) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int j = blockIdx.y;
    if (i < n && j < n && i < j) A[(size_t)j * ld + i] = 0.0;
    if (i == 0 && j == 0 && state[2] != 0) atomicCAS(info, 0, -1);
}

}  // namespace

// ---- host side ----
int chol_plan_tiles(int n) { return (n + TB_ - 1) / TB_; }
size_t chol_plan_state_ints(int n) {
    const size_t T = (size_t)chol_plan_tiles(n);
    return 4 + T * T;
}

// The bulk list. Step k's segment: the panel solves, then the next panel column, then the rest by columns. Every dependency of a task is
// produced by the critical CTA, by a priority CTA (whose own inputs come from the critical CTA and from earlier segments), or sits earlier in the list.
std::vector<uint32_t> chol_plan_tasks(int n) {
    const int T = chol_plan_tiles(n);
    std::vector<uint32_t> t;
    for (int k = 0; k + 1 < T; ++k) {
        // tile (k+1, k) and (k+1, k+1) of step k belong to the critical CTA
        // S(k+2, k), U(k+2, k+1, k), U(k+2, k+2, k) belong to the two priority CTAs (run_worker)
        for (int i = k + 3; i < T; ++i) t.push_back(pack_task(TASK_S, i, k, k));
        for (int i = k + 3; i < T; ++i) t.push_back(pack_task(TASK_U, i, k + 1, k));
        for (int j = k + 2; j < T; ++j)
            for (int i = j; i < T; ++i) {
                if (i == k + 2 && j == k + 2) continue;
                t.push_back(pack_task(TASK_U, i, j, k));
            }
    }
    return t;
}

cudaError_t chol_lower_persistent(cudaStream_t st, double* A, int ld, int n, int* info_dev, int* state_dev, const uint32_t* tasks_dev, int n_tasks,
                                  int max_ctas, long long* trace_dev) {
    const int T = chol_plan_tiles(n);
    if (T > 1023) return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(state_dev, 0, chol_plan_state_ints(n) * sizeof(int), st);
    if (e != cudaSuccess) return e;
    const size_t smem = sizeof(Shared) + 128;
    static const int minb = getenv("GPIRT_CHOL_MINB") ? atoi(getenv("GPIRT_CHOL_MINB")) : 1;
    auto kern = minb <= 1 ? k_chol_persistent<1> : (minb == 2 ? k_chol_persistent<2> : k_chol_persistent<3>);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int grid = 1;  // T <= 2: the critical CTA does everything
    if (T >= 3) {
        int bulk = n_tasks < max_ctas - 3 ? n_tasks : max_ctas - 3;
        if (n_tasks > 0 && bulk < 1) bulk = 1;  // the bulk list needs at least one CTA of its own
        grid = 3 + (bulk > 0 ? bulk : 0);
    }
    kern<<<grid, CP_THREADS, smem, st>>>(A, ld, n, T, state_dev, info_dev, tasks_dev, n_tasks, trace_dev);
    ++g_kernel_launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    dim3 zg((n + 255) / 256, n);
    k_zero_upper2<<<zg, 256, 0, st>>>(A, ld, n, state_dev, info_dev);
    ++g_kernel_launches;
    return cudaGetLastError();
}

}  // namespace gpirt

// ------------------------------------------------------------------------------------------------------------------------------------------
// gpirt_chol_device_noalloc (include/gpirt_b200.h): the persistent factorisation for callers that cannot allocate device memory — code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc. No counterpart in the reference.
// The kernel's scratch (tile version counters and the ordered task list, which the handle-based entry points keep in the sampler's
// arena or in temporaries of the call) lives in ONE file-scope device array per GPU that has loaded the library, held until the process exits.
// Every call on a device shares it: the state words are cleared and the task list is rewritten at the start of each call, and calls that overlap
// on the same GPU (two streams, two host threads) race on it — issue them on one stream, or serialise them.
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "../../include/gpirt_b200.h"

namespace gpirt { int fail_msg(int code, const char* msg); }  // capi.cu: sets gpirt_last_error()

namespace gpirt {
namespace {

constexpr int NOALLOC_MAX_N = GPIRT_NOALLOC_CHOL_MAX_N;                    // 2048 -> 32 x 32 tiles
constexpr int NOALLOC_MAX_T = (NOALLOC_MAX_N + TB_ - 1) / TB_;
constexpr size_t NOALLOC_STATE_BYTES = (((size_t)(4 + NOALLOC_MAX_T * NOALLOC_MAX_T) * sizeof(int)) + 255) & ~(size_t)255;
constexpr size_t NOALLOC_MAX_TASKS = 6144;                                  // the task list of n = 2048 has 5800 entries (checked at run time)
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_STATE_BYTES + NOALLOC_MAX_TASKS * sizeof(uint32_t);  // 28 928 bytes

// In HOST code this name is a stand-in, not a device address: the address the kernels get comes from cudaGetSymbolAddress, once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// number of bulk-list tasks of step k for T tiles per side (the order of chol_plan_tasks)
__host__ __device__ inline int tasks_of_step(int T, int k) {
    if (k + 1 >= T) return 0;
    const int s = T - k - 3 > 0 ? T - k - 3 : 0;   // S(i, k, k) and U(i, k+1, k), i >= k + 3
    const int r = T - k - 2;                        // columns j >= k + 2: r (r + 1) / 2 tiles, minus (k+2, k+2)
    return 2 * s + (r > 0 ? r * (r + 1) / 2 - 1 : 0);
}

// the task list of chol_plan_tasks, written on the device: thread k writes the segment of step k
__global__ void k_noalloc_plan(uint32_t* tasks, int T) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k + 1 >= T) return;
    int off = 0;
    for (int q = 0; q < k; ++q) off += tasks_of_step(T, q);
    uint32_t* t = tasks + off;
    for (int i = k + 3; i < T; ++i) *t++ = pack_task(TASK_S, i, k, k);
    for (int i = k + 3; i < T; ++i) *t++ = pack_task(TASK_U, i, k + 1, k);
    for (int j = k + 2; j < T; ++j)
        for (int i = j; i < T; ++i) {
            if (i == k + 2 && j == k + 2) continue;
            *t++ = pack_task(TASK_U, i, j, k);
        }
}

}  // namespace
}  // namespace gpirt

extern "C" int gpirt_chol_device_noalloc(int device, double* A_dev, int64_t ld, int64_t n, int32_t* info_dev, void* cuda_stream) {
    using namespace gpirt;
    char msg[256];
    if (!A_dev || !info_dev) return fail_msg(GPIRT_ERR_INVALID, "chol_device_noalloc: NULL device pointer");
    if (n < 1 || n > NOALLOC_MAX_N) {
        snprintf(msg, sizeof(msg), "chol_device_noalloc: n = %lld outside 1..%d (the static scratch of this entry point is sized for %d; use gpirt_chol or a sampler handle beyond it)",
                 (long long)n, NOALLOC_MAX_N, NOALLOC_MAX_N);
        return fail_msg(GPIRT_ERR_SHAPE, msg);
    }
    if (ld < ((n + 1) & ~(int64_t)1) || (ld & 1) || ((uintptr_t)A_dev & 15)) {
        snprintf(msg, sizeof(msg), "chol_device_noalloc: ld = %lld must be even and >= n rounded up to even (%lld), A 16-byte aligned (columns are moved by 16-byte bulk copies)",
                 (long long)ld, (long long)((n + 1) & ~(int64_t)1));
        return fail_msg(GPIRT_ERR_SHAPE, msg);
    }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return fail_msg(GPIRT_ERR_NO_DEVICE, "no CUDA device; gpirt_b200 has no CPU path");
    if (device < 0 || device >= count || device >= NOALLOC_MAX_DEVICES) return fail_msg(GPIRT_ERR_INVALID, "chol_device_noalloc: device out of range");
    cudaError_t e = cudaSetDevice(device);
    unsigned char* scratch = nullptr;
    if (e == cudaSuccess) {  // set-up, once per device, with that device current: the device address behind the host-side stand-in
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e == cudaSuccess) g_noalloc_ptr[device] = (unsigned char*)p;
        }
        scratch = g_noalloc_ptr[device];
    }
    if (e != cudaSuccess || !scratch) {
        snprintf(msg, sizeof(msg), "chol_device_noalloc: cudaGetSymbolAddress / cudaSetDevice: %s", cudaGetErrorString(e));
        return fail_msg(GPIRT_ERR_CUDA, msg);
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int T = chol_plan_tiles((int)n);
    int n_tasks = 0;
    for (int k = 0; k < T; ++k) n_tasks += tasks_of_step(T, k);
    if ((size_t)n_tasks > NOALLOC_MAX_TASKS) return fail_msg(GPIRT_ERR_SHAPE, "chol_device_noalloc: task list exceeds the static scratch");
    int* state = (int*)scratch;
    uint32_t* tasks = (uint32_t*)(scratch + NOALLOC_STATE_BYTES);
    // what must start at zero: the caller's info word here, the state words inside chol_lower_persistent; the task list is rewritten for this n
    if ((e = cudaMemsetAsync(info_dev, 0, sizeof(int32_t), st)) == cudaSuccess && T > 1) {
        k_noalloc_plan<<<1, 64, 0, st>>>(tasks, T);
        ++g_kernel_launches;
        e = cudaGetLastError();
    }
    static const int ctas = getenv("GPIRT_CHOL_CTAS") ? atoi(getenv("GPIRT_CHOL_CTAS")) : 48;
    if (e == cudaSuccess) e = chol_lower_persistent(st, A_dev, (int)ld, (int)n, (int*)info_dev, state, tasks, n_tasks, ctas);
    if (e != cudaSuccess) {
        snprintf(msg, sizeof(msg), "chol_device_noalloc: %s", cudaGetErrorString(e));
        return fail_msg(GPIRT_ERR_CUDA, msg);
    }
    return GPIRT_OK;
}
