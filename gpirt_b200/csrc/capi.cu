// C-ABI of the B200-native gpirt sweep: lifetime, data/state marshalling and the subsystem entry points
// declared in include/gpirt_b200.h (each declaration there cites the reference interface it replaces).
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <type_traits>
#include <vector>
#include <mutex>
#include <map>

#include "probit.cuh"
#include "sampler_state.cuh"

using namespace gpirt;

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
namespace gpirt { int fail_msg(int code, const char* msg) { return fail(code, "%s", msg); } }  // for entry.cu
#define CU(x)                                                                                           \
    do {                                                                                                \
        cudaError_t e_ = (x);                                                                           \
        if (e_ != cudaSuccess) return fail(GPIRT_ERR_CUDA, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)
#define CHECK_H(h) \
    if (!(h)) return fail(GPIRT_ERR_INVALID, "NULL sampler handle")
#define TRY_(x)                  \
    do {                         \
        int rc_ = (x);           \
        if (rc_ != GPIRT_OK) return rc_; \
    } while (0)

// Device allocations go through a small per-process cache: gpirtMCMC() creates and destroys a handle per call (arena, theta-step table, draw
// store, upload staging: ~1 GB at C3), and cudaMalloc / cudaFree of blocks that size map and unmap device memory — tens of milliseconds on a
// busy host, more than the sweeps of a short call. A freed block of >= 1 MB is kept for an exact-size reuse on the same device. The cache is
// bounded by GPIRT_CACHE_MB (default 4096 MB per process, 0 disables it; at most 32 blocks, oldest evicted) because other allocators in the
// process (torch, NCCL) cannot see or reclaim what is parked here; gpirt_release_cached_memory() empties it. Like cudaFree, dev_free
// synchronises the block's device first.
namespace {
struct CachedBlock { void* p; size_t bytes; int device; };
std::mutex g_alloc_mu;
std::vector<CachedBlock> g_free_blocks;
std::map<void*, std::pair<size_t, int>> g_live_blocks;
size_t g_cached_bytes = 0;
constexpr size_t CACHE_MIN_BLOCK = (size_t)1 << 20, CACHE_MAX_BLOCKS = 32;
size_t cache_max_bytes() {
    static const size_t cap = [] {
        const char* e = getenv("GPIRT_CACHE_MB");
        const long long mb = e ? atoll(e) : 4096;
        return (size_t)(mb > 0 ? mb : 0) << 20;
    }();
    return cap;
}
struct DeviceScope {  // run a few calls on another device, then return to the caller's
    int prev = -1;
    explicit DeviceScope(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != device) cudaSetDevice(device); else prev = -1;
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
};
cudaError_t free_on(void* p, int device) {
    DeviceScope ds(device);
    return cudaFree(p);
}
}  // namespace

static cudaError_t dev_malloc(void** p, size_t bytes) {
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return e;
    {
        std::lock_guard<std::mutex> lk(g_alloc_mu);
        for (size_t k = 0; k < g_free_blocks.size(); ++k)
            if (g_free_blocks[k].bytes == bytes && g_free_blocks[k].device == device) {
                *p = g_free_blocks[k].p;
                g_cached_bytes -= bytes;
                g_free_blocks.erase(g_free_blocks.begin() + k);
                g_live_blocks[*p] = {bytes, device};
                return cudaSuccess;
            }
    }
    e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) {  // out of memory with blocks parked in the cache: release them and retry once
        cudaGetLastError();
        std::vector<CachedBlock> drop;
        {
            std::lock_guard<std::mutex> lk(g_alloc_mu);
            drop.swap(g_free_blocks);
            g_cached_bytes = 0;
        }
        for (const CachedBlock& b : drop) free_on(b.p, b.device);
        e = cudaMalloc(p, bytes);
        if (e != cudaSuccess) return e;
    }
    std::lock_guard<std::mutex> lk(g_alloc_mu);
    g_live_blocks[*p] = {bytes, device};
    return cudaSuccess;
}

static cudaError_t dev_free(void* p) {
    if (!p) return cudaSuccess;
    size_t bytes = 0;
    int device = 0;
    {
        std::lock_guard<std::mutex> lk(g_alloc_mu);
        auto it = g_live_blocks.find(p);
        if (it == g_live_blocks.end()) return cudaFree(p);
        bytes = it->second.first;
        device = it->second.second;
        g_live_blocks.erase(it);
    }
    DeviceScope ds(device);  // the block's device, not whatever the calling thread has current
    if (bytes < CACHE_MIN_BLOCK || bytes > cache_max_bytes()) return cudaFree(p);
    cudaDeviceSynchronize();  // cudaFree's implicit synchronisation: nothing in flight may still use the block when it is handed out again
    std::vector<CachedBlock> evict;
    {
        std::lock_guard<std::mutex> lk(g_alloc_mu);
        g_free_blocks.push_back({p, bytes, device});
        g_cached_bytes += bytes;
        while (g_free_blocks.size() > CACHE_MAX_BLOCKS || g_cached_bytes > cache_max_bytes()) {
            evict.push_back(g_free_blocks.front());
            g_cached_bytes -= g_free_blocks.front().bytes;
            g_free_blocks.erase(g_free_blocks.begin());
        }
    }
    cudaError_t e = cudaSuccess;
    for (const CachedBlock& b : evict) {
        cudaError_t eb = free_on(b.p, b.device);
        if (eb != cudaSuccess) e = eb;
    }
    return e;
}

// temporaries of one call: freed on every return path
struct DevTemp {
    void* p = nullptr;
    DevTemp() = default;
    DevTemp(const DevTemp&) = delete;
    DevTemp& operator=(const DevTemp&) = delete;
    ~DevTemp() { if (p) dev_free(p); }
    cudaError_t alloc(size_t bytes) { return dev_malloc(&p, bytes); }
    template <class T> T* as() const { return (T*)p; }
};

extern "C" int gpirt_release_cached_memory(void) {
    std::vector<CachedBlock> drop;
    {
        std::lock_guard<std::mutex> lk(g_alloc_mu);
        drop.swap(g_free_blocks);
        g_cached_bytes = 0;
    }
    for (const CachedBlock& b : drop) free_on(b.p, b.device);
    return GPIRT_OK;
}

static int select_device(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(GPIRT_ERR_NO_DEVICE, "no CUDA device (%s); gpirt_b200 has no CPU path", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(GPIRT_ERR_INVALID, "device %d out of range (0..%d)", device, count - 1);
    CU(cudaSetDevice(device));
    return GPIRT_OK;
}

// SEC(h, "name", launch): times `launch` with CUDA events when profiling is on (never in the timed bench pass)
struct SecScope {
    gpirt_sampler* h; bool on; gpirt_sampler::Pending p; cudaStream_t st;
    SecScope(gpirt_sampler* h_, const char* name, cudaStream_t st_) : h(h_), on(h_->profiling), st(st_) {
        if (!on) return;
        p.name = name;
        cudaEventCreate(&p.a); cudaEventCreate(&p.b);
        cudaEventRecord(p.a, st);
    }
    ~SecScope() {
        if (!on) return;
        cudaEventRecord(p.b, st);
        h->pending.push_back(p);
    }
};
#define SEC(h, name, call) do { SecScope sec_(h, name, (h)->stream); call; } while (0)
#define SEC_ST(h, st, name, call) do { SecScope sec_(h, name, st); call; } while (0)

// main stream joins the side stream's work (no-ops when nothing is in flight)
static int wait_L(gpirt_sampler* h) {
    if (h->pending_L) {
        SEC(h, "wait_L", CU(cudaStreamWaitEvent(h->stream, h->ev_L, 0)));
        h->pending_L = false;
    }
    return GPIRT_OK;
}
static int wait_W(gpirt_sampler* h) {
    if (h->pending_W) {
        SEC(h, "wait_W", CU(cudaStreamWaitEvent(h->stream, h->ev_W, 0)));  // recorded after ev_L on the same stream
        h->pending_W = false;
        h->pending_L = false;
    }
    return GPIRT_OK;
}
static int wait_shared(gpirt_sampler* h) {
    if (h->pending_shared) {
        SEC(h, "wait_shared", CU(cudaStreamWaitEvent(h->stream, h->ev_shared, 0)));  // recorded after ev_L and ev_W on the same stream
        h->pending_shared = false;
        h->pending_L = false;
        h->pending_W = false;
    }
    return GPIRT_OK;
}
static int quiesce(gpirt_sampler* h) {
    TRY_(wait_L(h));
    return wait_shared(h);
}

#define TRY(x)                   \
    do {                         \
        int rc_ = (x);           \
        if (rc_ != GPIRT_OK) return rc_; \
    } while (0)

__global__ void k_add_diag(double* A, int ld, int n, double v) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) A[(size_t)i * ld + i] = __dadd_rn(A[(size_t)i * ld + i], v);
}
__global__ void k_fill_grid(double* t) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < NGRID) t[i] = grid_value(i);
}
// beta0[q,J] = mean + sd * z  (R::rnorm(mean, sd), src/gpirtMCMC.cpp:153)
__global__ void k_init_beta(double* beta, const double* means, const double* sds, int m_total, RngKey key) {
    int J = blockIdx.x * blockDim.x + threadIdx.x;
    if (J >= m_total) return;
    for (int q = 0; q < 3; ++q) beta[3 * (size_t)J + q] = means[3 * (size_t)J + q] + sds[3 * (size_t)J + q] * philox_normal(key, ST_INIT_BETA, J, q);
}

// thresholds item-major (C+1) x m -> the reference's m x (C+1) column-major slice of threshold_draws
// One kept sample in one launch: theta, beta, the thresholds (m x (C+1) out of the (C+1) x m state) and the total log-likelihood — the last CTA adds the
// per-item terms as k_ordered_sum does (256 strided partial sums combined in index order by one thread: the same bits).
__global__ void __launch_bounds__(256) k_store_sample(const double* __restrict__ theta, int n, const double* __restrict__ beta, int m,
                                                      const double* __restrict__ thr, int C1, const double* __restrict__ item_ll,
                                                      double* __restrict__ o_theta, double* __restrict__ o_beta, double* __restrict__ o_thr,
                                                      double* __restrict__ o_ll) {
    if (blockIdx.x == gridDim.x - 1) {
        __shared__ double part[256];
        double acc = 0.0;
        for (int i = threadIdx.x; i < m; i += 256) acc += item_ll[i];
        part[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int k = 0; k < 256; ++k) t += part[k];
            *o_ll = t;
        }
        return;
    }
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < n) o_theta[e] = theta[e];
    if (e < 3 * m) o_beta[e] = beta[e];
    if (e < C1 * m) {
        const int c = e / m, J = e - c * m;  // e = c * m + J
        o_thr[e] = thr[(size_t)J * C1 + c];
    }
}

extern "C" {

const char* gpirt_last_error(void) { return g_err; }

int gpirt_device_count(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) return 0;
    return count;
}

int gpirt_delta_to_threshold(const double* d, int C, double* thr) {  // src/log-likelihood.cpp:66-76
    if (!d || !thr || C < 2) return fail(GPIRT_ERR_INVALID, "delta_to_threshold: bad arguments");
    for (int i = 0; i <= C; ++i) thr[i] = 0.0;
    thr[0] = -std::numeric_limits<double>::infinity();
    thr[1] = d[0];
    thr[C] = std::numeric_limits<double>::infinity();
    for (int i = 1; i < C - 1; ++i) thr[i + 1] = thr[i] + std::exp(d[i]);
    return GPIRT_OK;
}
int gpirt_threshold_to_delta(const double* thr, int C, double* d) {  // src/log-likelihood.cpp:78-88
    if (!d || !thr || C < 2) return fail(GPIRT_ERR_INVALID, "threshold_to_delta: bad arguments");
    for (int i = 0; i < C - 1; ++i) d[i] = 0.0;
    d[0] = thr[1];
    for (int i = 1; i < C - 1; ++i) d[i] = std::log(thr[i + 1] - thr[i]);
    return GPIRT_OK;
}

int gpirt_create(const gpirt_config* cfg, gpirt_sampler** out) {
    if (!cfg || !out) return fail(GPIRT_ERR_INVALID, "gpirt_create: NULL argument");
    *out = nullptr;
    if (cfg->horizon != 1) return fail(GPIRT_ERR_UNSUPPORTED, "horizon %d: only horizon == 1 is built", cfg->horizon);
    if (cfg->constant_IRF != 0) return fail(GPIRT_ERR_UNSUPPORTED, "constant_IRF %d: only 0 is built", cfg->constant_IRF);
    if (cfg->C < 2 || cfg->C > GPIRT_MAX_C) return fail(GPIRT_ERR_CATEGORIES, "C = %d outside 2..%d", cfg->C, GPIRT_MAX_C);
    if (cfg->n < 1 || cfg->m_total < 1) return fail(GPIRT_ERR_SHAPE, "n = %d, m = %d", cfg->n, cfg->m_total);
    if (cfg->p_beta != 2 && cfg->p_beta != 3) return fail(GPIRT_ERR_SHAPE, "p_beta = %d (2 or 3)", cfg->p_beta);
    if (cfg->world < 1 || cfg->rank < 0 || cfg->rank >= cfg->world) return fail(GPIRT_ERR_INVALID, "rank %d of %d", cfg->rank, cfg->world);
    if (cfg->m_total < cfg->world || cfg->n < cfg->world) return fail(GPIRT_ERR_SHAPE, "fewer items/respondents than ranks");
    if (cfg->kernel != 0 && cfg->kernel != 1) return fail(GPIRT_ERR_INVALID, "KERNEL must be Matern (0) or RBF (1)");
    TRY(select_device(cfg->device));

    gpirt_sampler* h = new gpirt_sampler();
    h->cfg = *cfg;
    h->n = cfg->n; h->m_total = cfg->m_total; h->C = cfg->C; h->world = cfg->world; h->rank = cfg->rank;
    h->m_chunk = (h->m_total + h->world - 1) / h->world;
    h->item0 = h->rank * h->m_chunk;
    h->m_local = std::max(0, std::min(h->m_chunk, h->m_total - h->item0));
    h->n_chunk = (h->n + h->world - 1) / h->world;
    h->resp0 = h->rank * h->n_chunk;
    h->n_local = std::max(0, std::min(h->n_chunk, h->n - h->resp0));
    h->ldn = pad_ld(h->n); h->ldN = pad_ld(NGRID); h->ldy = pad_ld(h->n);
    h->ld_chunk = pad_ld(h->m_chunk); h->ldz = std::max(h->ldn, h->ldN);
    h->ldm = pad_ld((long long)h->world * h->m_chunk);
    h->ldml = pad_ld(h->m_chunk);
    // regime switch of src/draw-theta.cpp:112,171,206 at horizon 1: CST and RDM share L_i = sqrt(1 + sd^2)
    const double ls = cfg->theta_ls;
    h->theta_regime = (ls >= 3.0 || ls <= 0.1) ? 0 : 1;
    *out = h;  // from here on the caller can destroy a half-built handle

    CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    h->own_stream = true;
    {
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU(cudaStreamCreateWithPriority(&h->side, cudaStreamNonBlocking, hi));
        CU(cudaStreamCreateWithPriority(&h->side2, cudaStreamNonBlocking, hi));
        CU(cudaEventCreateWithFlags(&h->ev_tmp, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_L, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_shared, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_W, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_xchg, cudaEventDisableTiming));
        if (h->n >= 4096) {  // look-ahead of the two-level Cholesky (linalg.cuh, CholAux): the bulk of a trailing update runs beside the next panels
            CU(cudaStreamCreateWithPriority(&h->chol_aux.helper, cudaStreamNonBlocking, lo));  // below the side stream: the panels must not queue behind the bulk GEMM's pending tiles
            CU(cudaEventCreateWithFlags(&h->chol_aux.ev_panel, cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->chol_aux.ev_rest, cudaEventDisableTiming));
        }
        // milestones of chol(K): rows 50 %, 75 %, 87.5 %, 93.75 % and 100 % (multiples of the panel width), see do_draw_f
        int prev = 0;
        for (double frac : {0.5, 0.75, 0.875, 0.9375, 1.0}) {
            int r = frac >= 1.0 ? h->n : ((int)(h->n * frac) / CHOL_NB) * CHOL_NB;
            if (r <= prev) continue;
            CU(cudaEventCreateWithFlags(&h->L_rows.ev[h->L_rows.count], cudaEventDisableTiming));
            h->L_rows.rows[h->L_rows.count++] = r;
            prev = r;
        }
        prev = 0;
        for (double frac : {0.5, 0.75, 0.875, 0.9375, 1.0}) {
            int r = frac >= 1.0 ? (int)NGRID : ((int)(NGRID * frac) / CHOL_NB) * CHOL_NB;
            if (r <= prev) continue;
            CU(cudaEventCreateWithFlags(&h->Lpost_rows.ev[h->Lpost_rows.count], cudaEventDisableTiming));
            h->Lpost_rows.rows[h->Lpost_rows.count++] = r;
            prev = r;
        }
    }
    const size_t mt = (size_t)h->world * h->m_chunk;
    // one device allocation for everything a handle owns for life (create + destroy cost one cudaMalloc / cudaFree, which is what an
    // end-to-end gpirtMCMC() call of a few sweeps notices), carved into 256-byte aligned pieces, zero-filled on the handle's stream
    struct Piece { void** p; size_t bytes; };
    std::vector<Piece> pieces;
    size_t arena_bytes = 0;
    auto want = [&](auto** p, size_t count) {
        using T = std::remove_pointer_t<std::remove_pointer_t<decltype(p)>>;
        const size_t b = ((std::max<size_t>(count, 1) * sizeof(T)) + 255) & ~(size_t)255;
        pieces.push_back({(void**)p, b});
        arena_bytes += b;
    };
    want(&h->y_im, (size_t)h->ldy * std::max(h->m_local, 1));
    want(&h->y_rm, (size_t)h->ldm * std::max(h->n_local, 1));
    if (h->world > 1) {
        want(&h->y_rl, (size_t)h->ldml * h->n);
        want(&h->theta_part, (size_t)h->world * h->n * (THETA_BATCH + 1));
        want(&h->theta_ll0, (size_t)h->n);
        want(&h->theta_knext, (size_t)h->n);
        want(&h->theta_nact, 1);
        want(&h->items_packed, (size_t)h->world * h->m_chunk * (h->C + 5));
    }
    want(&h->theta, (size_t)h->world * h->n_chunk);
    want(&h->theta_x, (size_t)h->world * h->n_chunk);
    want(&h->beta, 3 * mt);
    want(&h->thr, (size_t)(h->C + 1) * mt);
    want(&h->f, (size_t)h->ldn * h->m_local);
    want(&h->nu, (size_t)h->ldn * h->m_local);
    want(&h->zbuf, (size_t)h->ldz * h->m_local);
    want(&h->fstar, (size_t)h->world * NGRID * h->ld_chunk);
    want(&h->L, (size_t)h->ldn * h->n);
    want(&h->W, (size_t)h->ldn * NGRID);
    want(&h->Kpost, (size_t)h->ldN * NGRID);
    want(&h->tstar, (size_t)h->ldN);
    want(&h->sds, 3 * mt);
    want(&h->tps, 2 * (size_t)h->n);
    want(&h->item_ll, mt);
    want(&h->total_ll, 1);
    want(&h->flags, 1);
    want(&h->flags_x, gpirt::MAX_PEERS + 1);
    want(&h->flags_c, gpirt::MAX_PEERS + 1);
    want(&h->chol_ws, chol_ws_ints(std::max(h->n, (int)NGRID)));
    const std::vector<uint32_t> tasks_n = h->n < 4096 ? chol_plan_tasks(h->n) : std::vector<uint32_t>(), tasks_N = chol_plan_tasks(NGRID);
    want(&h->chol_tasks_n, tasks_n.size());
    want(&h->chol_tasks_N, tasks_N.size());
    want(&h->chol_state_n, chol_plan_state_ints(h->n < 4096 ? h->n : 64));
    want(&h->chol_state_N, chol_plan_state_ints(NGRID));
    want(&h->Linv, (size_t)h->ldn * h->n);  // block-inverse solver of draw_fstar (the default)
    want(&h->W2, (size_t)h->ldn * NGRID);
    want(&h->tri_scratch, trtri_scratch_elems(h->n));
    for (int b = 0; b < 5; ++b) {
        size_t units = (b == GPIRT_BLOCK_THETA) ? (size_t)h->world * h->n_chunk : mt;
        want(&h->evals[b], units);
        want(&h->exact[b], units);
        want(&h->margin[b], units);
    }
    CU(dev_malloc((void**)&h->arena, arena_bytes));
    CU(cudaMemsetAsync(h->arena, 0, arena_bytes, h->stream));
    {
        size_t off = 0;
        for (const Piece& pc : pieces) { *pc.p = h->arena + off; off += pc.bytes; }
    }
    {
        static const int chol_ctas = getenv("GPIRT_CHOL_CTAS") ? atoi(getenv("GPIRT_CHOL_CTAS")) : 48;
        if (!tasks_n.empty()) CU(cudaMemcpyAsync(h->chol_tasks_n, tasks_n.data(), tasks_n.size() * 4, cudaMemcpyHostToDevice, h->stream));
        if (!tasks_N.empty()) CU(cudaMemcpyAsync(h->chol_tasks_N, tasks_N.data(), tasks_N.size() * 4, cudaMemcpyHostToDevice, h->stream));
        CU(cudaStreamSynchronize(h->stream));  // the host vectors go out of scope
        if (h->n < 4096) {
            h->plan_n.tasks = h->chol_tasks_n; h->plan_n.n_tasks = (int)tasks_n.size(); h->plan_n.state = h->chol_state_n; h->plan_n.max_ctas = chol_ctas;
        }
        h->plan_N.tasks = h->chol_tasks_N; h->plan_N.n_tasks = (int)tasks_N.size(); h->plan_N.state = h->chol_state_N; h->plan_N.max_ctas = chol_ctas;
    }
    k_fill_grid<<<(NGRID + 255) / 256, 256, 0, h->stream>>>(h->tstar);
    ++g_kernel_launches;
    CU(cudaGetLastError());
    return GPIRT_OK;
}

static void store_free(gpirt_sampler* h);

int gpirt_destroy(gpirt_sampler* h) {
    if (!h) return GPIRT_OK;
    cudaSetDevice(h->cfg.device);
    if (h->side) cudaStreamSynchronize(h->side);
    if (h->side2) cudaStreamSynchronize(h->side2);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_L) cudaEventDestroy(h->ev_L);
    if (h->ev_shared) cudaEventDestroy(h->ev_shared);
    if (h->ev_W) cudaEventDestroy(h->ev_W);
    if (h->ev_xchg) cudaEventDestroy(h->ev_xchg);
    for (int c = 0; c < h->L_rows.count; ++c) cudaEventDestroy(h->L_rows.ev[c]);
    for (int c = 0; c < h->Lpost_rows.count; ++c) cudaEventDestroy(h->Lpost_rows.ev[c]);
    if (h->side) cudaStreamDestroy(h->side);
    if (h->side2) cudaStreamDestroy(h->side2);
    if (h->ev_tmp) cudaEventDestroy(h->ev_tmp);
    if (h->chol_aux.helper) { cudaStreamSynchronize(h->chol_aux.helper); cudaStreamDestroy(h->chol_aux.helper); }
    if (h->chol_aux.ev_panel) cudaEventDestroy(h->chol_aux.ev_panel);
    if (h->chol_aux.ev_rest) cudaEventDestroy(h->chol_aux.ev_rest);
    for (int p = 0; p <= gpirt::MAX_PEERS; ++p)
        if (h->chain_peer[p] && h->chain_peer_ipc[p] && h->chain_peer[p] != h->peer_arena[p]) cudaIpcCloseMemHandle(h->chain_peer[p]);
    for (char*& pa : h->peer_arena)
        if (pa) { cudaIpcCloseMemHandle(pa); pa = nullptr; }
    for (void* p : {(void*)h->d_z, (void*)h->d_u, (void*)h->theta_tab, (void*)h->mu_rows, (void*)h->arena})
        if (p) dev_free(p);
    store_free(h);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return GPIRT_OK;
}

int gpirt_set_stream(gpirt_sampler* h, void* cuda_stream) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    TRY(quiesce(h));
    CU(cudaStreamSynchronize(h->stream));
    if (h->own_stream) { cudaStreamDestroy(h->stream); h->own_stream = false; }
    h->stream = (cudaStream_t)cuda_stream;
    return GPIRT_OK;
}

int gpirt_set_solver(gpirt_sampler* h, int solver) {
    CHECK_H(h);
    if (solver != 0 && solver != 1) return fail(GPIRT_ERR_INVALID, "solver %d: 0 (blocked substitution) or 1 (block inverse)", solver);
    h->solver = solver;
    return GPIRT_OK;
}

// CTAs the persistent Cholesky kernels of this handle may hold (chol_persistent.cu); 0: the per-panel path of linalg.cu. Several chains side by side
// (gpirt_mcmc_chains) share the SMs: every chain's persistent kernel parks spinning CTAs, so the budget is divided among them.
int gpirt_set_chol_ctas(gpirt_sampler* h, int ctas) {
    CHECK_H(h);
    if (ctas < 0) return fail(GPIRT_ERR_INVALID, "set_chol_ctas: %d", ctas);
    TRY(quiesce(h));
    if (ctas == 0) {
        h->plan_n.state = nullptr;
        h->plan_N.state = nullptr;
    } else {
        if (h->n < 4096) h->plan_n.state = h->chol_state_n;
        h->plan_N.state = h->chol_state_N;
        h->plan_n.max_ctas = h->plan_N.max_ctas = ctas < 4 ? 4 : ctas;
    }
    return GPIRT_OK;
}

int gpirt_debug_theta_threads(int threads) {
    if (threads != 0 && threads != 256 && threads != 1024) return fail(GPIRT_ERR_INVALID, "theta threads %d: 0 (automatic), 256 or 1024", threads);
    g_theta_threads = threads;
    return GPIRT_OK;
}

int gpirt_set_screen(gpirt_sampler* h, int on) {  // 0: every evaluation in fp64; 1: fp32 screen (default); 2: fp32 screen + early rejection on partial sums
    CHECK_H(h);
    h->screen = on != 0;
    if (on == 2) h->early_exit = 1;
    return GPIRT_OK;
}
int gpirt_get_exact_evals(gpirt_sampler* h, int block, int32_t* out) {
    CHECK_H(h);
    if (block < 0 || block > 4 || !out) return fail(GPIRT_ERR_INVALID, "get_exact_evals: block %d", block);
    CU(cudaSetDevice(h->cfg.device));
    const size_t units = (block == GPIRT_BLOCK_THETA) ? h->n : h->m_total;
    CU(cudaMemcpyAsync(out, h->exact[block], units * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return GPIRT_OK;
}

int gpirt_set_overlap(gpirt_sampler* h, int on) {
    CHECK_H(h);
    TRY(quiesce(h));
    h->overlap = on != 0;
    return GPIRT_OK;
}

int gpirt_set_iter(gpirt_sampler* h, uint32_t iter) {
    CHECK_H(h);
    h->iter = iter;
    return GPIRT_OK;
}

}  // extern "C"

// ---- deferred error flags ------------------------------------------------------------------------
static int check_flags(gpirt_sampler* h) {
    DevFlags fl;
    TRY(quiesce(h));
    CU(cudaMemcpyAsync(&fl, h->flags, sizeof(fl), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (fl.chol_info || fl.chol_post_info || fl.overrun || fl.stalled || fl.bad_code || fl.peer_timeout) {
        CU(cudaMemsetAsync(h->flags, 0, sizeof(fl), h->stream));
        if (fl.peer_timeout) return fail(GPIRT_ERR_CUDA, "peer exchange (f* chunks / the distributed theta-only chain): what rank %d publishes did not arrive within 4 s", fl.peer_timeout - 1);
        if (fl.bad_code) return fail(GPIRT_ERR_CATEGORIES, "response code outside 1..%d", h->C);
        if (fl.chol_info < 0 || fl.chol_post_info < 0) return fail(GPIRT_ERR_CUDA, "chol(): the persistent factorisation kernel gave up waiting for a tile (device stalled?)");
        if (fl.chol_info) return fail(GPIRT_ERR_NOT_PD, "chol(): decomposition failed: K(theta,theta) + 1e-6 I not positive definite at leading minor %d", fl.chol_info);
        if (fl.chol_post_info) return fail(GPIRT_ERR_NOT_PD, "chol(): decomposition failed: K_post not positive definite at leading minor %d", fl.chol_post_info);
        if (fl.overrun) return fail(GPIRT_ERR_STREAM_OVERRUN, "a supplied uniform stream ran dry before its ESS loop accepted");
        return fail(GPIRT_ERR_ESS_STALLED, "an ESS loop made %d proposals without accepting (NaN likelihood?)", ESS_MAX_PROPOSALS);
    }
    return GPIRT_OK;
}

static int upload(gpirt_sampler* h, double** buf, size_t* cap, const double* host, size_t count) {
    if (count > *cap) {
        if (*buf) CU(dev_free(*buf));
        *buf = nullptr;
        CU(dev_malloc((void**)buf, count * sizeof(double)));
        *cap = count;
    }
    CU(cudaMemcpyAsync(*buf, host, count * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    return GPIRT_OK;
}

static int build_layouts(gpirt_sampler* h, const int8_t* d_full) {
    CU(cudaMemsetAsync(h->y_im, 0, (size_t)h->ldy * std::max(h->m_local, 1), h->stream));
    CU(cudaMemsetAsync(h->y_rm, 0, (size_t)h->ldm * std::max(h->n_local, 1), h->stream));
    CU(launch_y_layouts(h->stream, d_full, h->n, h->m_total, h->item0, h->m_local, h->y_im, h->ldy, h->resp0, h->n_local, h->y_rm, h->ldm));
    if (h->y_rl && h->m_local > 0) {  // all respondents' rows over the local items: the local columns as a matrix of their own, row layout only
        CU(cudaMemsetAsync(h->y_rl, 0, (size_t)h->ldml * h->n, h->stream));
        CU(launch_y_layouts(h->stream, d_full + (size_t)h->item0 * h->n, h->n, h->m_local, 0, 0, h->y_im, h->ldy, 0, h->n, h->y_rl, h->ldml));
    }
    h->have_data = true;
    return GPIRT_OK;
}

extern "C" {

int gpirt_sync(gpirt_sampler* h) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    return check_flags(h);
}

// Rows [row0, row0 + nrows) x columns [col0, col0 + ncols) of the host matrix y (n x m_total doubles, NaN = missing) as int8 codes on the
// device (nrows x ncols, column-major), staged in column slabs so the fp64 copy never needs more than 256 MB of device memory.
static int upload_block_i8(gpirt_sampler* h, cudaStream_t st, const double* y, long long row0, long long nrows, long long col0, long long ncols,
                           int8_t* d_out, double* d_stage, long long slab_cols) {
    for (long long c0 = 0; c0 < ncols; c0 += slab_cols) {
        const long long cols = std::min<long long>(slab_cols, ncols - c0);
        CU(cudaMemcpy2DAsync(d_stage, (size_t)nrows * 8, y + (col0 + c0) * h->n + row0, (size_t)h->n * 8, (size_t)nrows * 8, (size_t)cols,
                             cudaMemcpyHostToDevice, st));
        CU(launch_y_to_i8(st, d_stage, cols * nrows, h->C, d_out + c0 * nrows, h->flags));
    }
    return GPIRT_OK;
}

int gpirt_set_data_f64(gpirt_sampler* h, const double* y) {
    CHECK_H(h);
    if (!y) return fail(GPIRT_ERR_INVALID, "set_data: NULL y");
    CU(cudaSetDevice(h->cfg.device));
    // A shard uploads what it keeps: its items' columns (item-major layout) and its respondents' rows (theta step) — 2/world of y.
    struct Block { long long row0, nrows, col0, ncols; };
    std::vector<Block> blocks;
    if (h->world == 1) blocks.push_back({0, h->n, 0, h->m_total});
    else {
        blocks.push_back({0, h->n, h->item0, h->m_local});
        blocks.push_back({h->resp0, h->n_local, 0, h->m_total});
    }
    long long max_rows = 1, max_count = 1;
    for (const Block& b : blocks) { max_rows = std::max(max_rows, b.nrows); max_count = std::max(max_count, b.nrows * b.ncols); }
    const long long slab_cols = std::max(1LL, (256LL << 20) / (8LL * max_rows));
    DevTemp t_y, t_i8;
    CU(t_i8.alloc((size_t)max_count));
    CU(t_y.alloc((size_t)std::min<long long>(slab_cols * max_rows, max_count) * 8));
    double* d_y = t_y.as<double>();
    int8_t* d_i8 = t_i8.as<int8_t>();
    int rc = GPIRT_OK;
    // The first upload of a handle runs on the side stream: nothing queued on the main stream can be reading the responses yet, so the initial draws
    // (K, its Cholesky, f0 = L z: gpirt_mcmc issues them first) run under the 8 n m bytes coming over PCIe. Replacing data joins everything first.
    cudaStream_t st = h->stream;
    if (!h->have_data && h->side && !h->pending_L && !h->pending_shared) st = h->side;
    else { TRY(quiesce(h)); CU(cudaStreamSynchronize(h->stream)); }
    CU(cudaMemsetAsync(h->y_im, 0, (size_t)h->ldy * std::max(h->m_local, 1), st));
    CU(cudaMemsetAsync(h->y_rm, 0, (size_t)h->ldm * std::max(h->n_local, 1), st));
    if (h->y_rl) CU(cudaMemsetAsync(h->y_rl, 0, (size_t)h->ldml * h->n, st));
    for (size_t k = 0; k < blocks.size() && rc == GPIRT_OK; ++k) {
        const Block& b = blocks[k];
        if (b.nrows <= 0 || b.ncols <= 0) continue;
        rc = upload_block_i8(h, st, y, b.row0, b.nrows, b.col0, b.ncols, d_i8, d_y, slab_cols);
        if (rc != GPIRT_OK) break;
        cudaError_t e;
        if (h->world == 1)
            e = launch_y_layouts(st, d_i8, h->n, h->m_total, 0, h->m_total, h->y_im, h->ldy, 0, h->n, h->y_rm, h->ldm);
        else if (k == 0)  // the block IS the local items: item-major layout, and all respondents' rows over these items (item-sharded theta step)
            e = launch_y_layouts(st, d_i8, h->n, h->m_local, 0, h->m_local, h->y_im, h->ldy, 0, h->n, h->y_rl, h->ldml);
        else              // the block IS the local respondents: respondent-major layout only
            e = launch_y_layouts(st, d_i8, h->n_local, h->m_total, 0, 0, h->y_im, h->ldy, 0, h->n_local, h->y_rm, h->ldm);
        if (e != cudaSuccess) rc = fail(GPIRT_ERR_CUDA, "y layouts: %s", cudaGetErrorString(e));
    }
    cudaStreamSynchronize(st);  // the staging blocks go back to the cache below, and whatever the caller launches next comes after the layouts
    if (rc != GPIRT_OK) return rc;
    h->have_data = true;
    return check_flags(h);
}

int gpirt_set_data_i8(gpirt_sampler* h, const int8_t* y) {
    CHECK_H(h);
    if (!y) return fail(GPIRT_ERR_INVALID, "set_data: NULL y");
    CU(cudaSetDevice(h->cfg.device));
    const size_t count = (size_t)h->n * h->m_total;
    for (size_t e = 0; e < count; ++e)
        if (y[e] < 0 || y[e] > h->C) return fail(GPIRT_ERR_CATEGORIES, "response code %d outside 0..%d at element %zu", (int)y[e], h->C, e);
    DevTemp t_full;
    CU(t_full.alloc(count));
    CU(cudaMemcpyAsync(t_full.p, y, count, cudaMemcpyHostToDevice, h->stream));
    int rc = build_layouts(h, t_full.as<int8_t>());
    cudaStreamSynchronize(h->stream);
    return rc;
}

// Batched chains (BASELINE config 5, R/gpirtMCMC.R:111 `for(chain in 1:CHAIN)`): every chain scores the same responses, so chains after the
// first take the two int8 layouts from a handle that already holds them (device-to-device, 2 n m bytes) instead of re-uploading 8 n m bytes.
int gpirt_copy_data(gpirt_sampler* h, gpirt_sampler* src) {
    CHECK_H(h);
    CHECK_H(src);
    if (!src->have_data) return fail(GPIRT_ERR_INVALID, "copy_data: the source handle holds no response data");
    if (h->cfg.device != src->cfg.device || h->n != src->n || h->m_total != src->m_total || h->C != src->C || h->world != src->world ||
        h->rank != src->rank || h->ldy != src->ldy || h->ldm != src->ldm)
        return fail(GPIRT_ERR_SHAPE, "copy_data: the two handles differ in shape, shard or device");
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaStreamSynchronize(src->stream));
    CU(cudaMemcpyAsync(h->y_im, src->y_im, (size_t)h->ldy * std::max(h->m_local, 1), cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaMemcpyAsync(h->y_rm, src->y_rm, (size_t)h->ldm * std::max(h->n_local, 1), cudaMemcpyDeviceToDevice, h->stream));
    if (h->y_rl && src->y_rl) CU(cudaMemcpyAsync(h->y_rl, src->y_rl, (size_t)h->ldml * h->n, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->have_data = true;
    return GPIRT_OK;
}

// The stacked sampler of a constant-IRF panel (src/draw-f.cpp:84-139, src/draw_threshold.cpp:5-96): its n = n_resp * horizon rows belong to
// different sessions, and row (h, i) has the mean X_h(i,:) beta_h of ITS session — not a function of one beta per item. mu: n x m_total host
// matrix used by draw_f and draw_threshold instead of X beta; NULL returns to X beta. world == 1 only.
int gpirt_set_mean_rows(gpirt_sampler* h, const double* mu) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    if (h->world != 1) return fail(GPIRT_ERR_UNSUPPORTED, "set_mean_rows on an item shard");
    TRY(quiesce(h));
    CU(cudaStreamSynchronize(h->stream));
    if (!mu) {
        if (h->mu_rows) { dev_free(h->mu_rows); h->mu_rows = nullptr; }
        return GPIRT_OK;
    }
    if (!h->mu_rows) {
        CU(dev_malloc((void**)&h->mu_rows, (size_t)h->ldn * std::max(h->m_local, 1) * 8));
        CU(cudaMemsetAsync(h->mu_rows, 0, (size_t)h->ldn * std::max(h->m_local, 1) * 8, h->stream));
    }
    CU(cudaMemcpy2DAsync(h->mu_rows, (size_t)h->ldn * 8, mu, (size_t)h->n * 8, (size_t)h->n * 8, h->m_local, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return GPIRT_OK;
}

int gpirt_set_priors(gpirt_sampler* h, const double* beta_prior_sds, const double* theta_prior_sds) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    if (beta_prior_sds) {
        CU(cudaMemcpyAsync(h->sds, beta_prior_sds, 3 * (size_t)h->m_total * 8, cudaMemcpyHostToDevice, h->stream));
        for (int q = 0; q < 3; ++q) h->sds0[q] = beta_prior_sds[q];  // column 0 feeds every K() call
        TRY(quiesce(h));
        h->chol_valid = h->shared_valid = false;
    }
    if (theta_prior_sds) CU(cudaMemcpyAsync(h->tps, theta_prior_sds, 2 * (size_t)h->n * 8, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return GPIRT_OK;
}

int gpirt_set_state(gpirt_sampler* h, const double* theta, const double* beta, const double* f, const double* thresholds,
                    const double* fstar) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    cudaStream_t st = h->stream;
    if (theta) {
        TRY(quiesce(h));
        CU(cudaMemcpyAsync(h->theta, theta, (size_t)h->n * 8, cudaMemcpyHostToDevice, st));
        CU(launch_clamp_theta(st, h->theta, h->n));
        h->chol_valid = h->shared_valid = false;
    }
    if (beta) {  // p_beta x m_total -> 3 x m_total (third row zero when p_beta == 2)
        std::vector<double> b3(3 * (size_t)h->m_total, 0.0);
        for (int J = 0; J < h->m_total; ++J)
            for (int q = 0; q < h->cfg.p_beta; ++q) b3[3 * (size_t)J + q] = beta[(size_t)h->cfg.p_beta * J + q];
        CU(cudaMemcpyAsync(h->beta, b3.data(), b3.size() * 8, cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
    }
    if (thresholds) {  // m_total x (C+1) column-major -> item-major (C+1) x m_total
        const int C1 = h->C + 1;
        std::vector<double> t((size_t)C1 * h->m_total);
        for (int J = 0; J < h->m_total; ++J)
            for (int c = 0; c < C1; ++c) t[(size_t)J * C1 + c] = thresholds[(size_t)c * h->m_total + J];
        CU(cudaMemcpyAsync(h->thr, t.data(), t.size() * 8, cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
    }
    if (f && h->m_local > 0)
        CU(cudaMemcpy2DAsync(h->f, (size_t)h->ldn * 8, f + (size_t)h->item0 * h->n, (size_t)h->n * 8, (size_t)h->n * 8, h->m_local,
                             cudaMemcpyHostToDevice, st));
    if (fstar) {  // NGRID x m_total column-major -> every chunk of the grid-row-major table
        DevTemp t_tmp;
        CU(t_tmp.alloc((size_t)NGRID * h->m_total * 8));
        double* tmp = t_tmp.as<double>();
        CU(cudaMemcpyAsync(tmp, fstar, (size_t)NGRID * h->m_total * 8, cudaMemcpyHostToDevice, st));
        for (int r = 0; r < h->world; ++r) {
            const int j0 = r * h->m_chunk, cols = std::min(h->m_chunk, h->m_total - j0);
            if (cols > 0) CU(launch_transpose(st, tmp + (size_t)j0 * NGRID, NGRID, NGRID, cols, h->fstar + (size_t)r * NGRID * h->ld_chunk, h->ld_chunk));
        }
        CU(cudaStreamSynchronize(st));
    }
    CU(cudaStreamSynchronize(st));
    return GPIRT_OK;
}

int gpirt_get_state(gpirt_sampler* h, double* theta, double* beta, double* f, double* thresholds, double* fstar) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    TRY(check_flags(h));
    cudaStream_t st = h->stream;
    if (theta) CU(cudaMemcpyAsync(theta, h->theta, (size_t)h->n * 8, cudaMemcpyDeviceToHost, st));
    if (beta) {
        std::vector<double> b3(3 * (size_t)h->m_total);
        CU(cudaMemcpyAsync(b3.data(), h->beta, b3.size() * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (int J = 0; J < h->m_total; ++J)
            for (int q = 0; q < h->cfg.p_beta; ++q) beta[(size_t)h->cfg.p_beta * J + q] = b3[3 * (size_t)J + q];
    }
    if (thresholds) {
        const int C1 = h->C + 1;
        std::vector<double> t((size_t)C1 * h->m_total);
        CU(cudaMemcpyAsync(t.data(), h->thr, t.size() * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (int J = 0; J < h->m_total; ++J)
            for (int c = 0; c < C1; ++c) thresholds[(size_t)c * h->m_total + J] = t[(size_t)J * C1 + c];
    }
    if (f && h->m_local > 0)  // only the local items' columns are written (world > 1: the caller gathers)
        CU(cudaMemcpy2DAsync(f + (size_t)h->item0 * h->n, (size_t)h->n * 8, h->f, (size_t)h->ldn * 8, (size_t)h->n * 8, h->m_local,
                             cudaMemcpyDeviceToHost, st));
    if (fstar) {
        DevTemp t_tmp;
        CU(t_tmp.alloc((size_t)NGRID * h->m_total * 8));
        double* tmp = t_tmp.as<double>();
        for (int r = 0; r < h->world; ++r) {
            const int j0 = r * h->m_chunk, cols = std::min(h->m_chunk, h->m_total - j0);
            // chunk is cols x NGRID (ld_chunk) column-major in (j, idx); transpose back to NGRID x cols
            if (cols > 0) CU(launch_transpose(st, h->fstar + (size_t)r * NGRID * h->ld_chunk, h->ld_chunk, cols, NGRID, tmp + (size_t)j0 * NGRID, NGRID));
        }
        CU(cudaMemcpyAsync(fstar, tmp, (size_t)NGRID * h->m_total * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    CU(cudaStreamSynchronize(st));
    return GPIRT_OK;
}


// ---- device-side draw store ----
static void store_free(gpirt_sampler* h) {
    auto& s = h->store;
    for (double* p : {s.theta, s.beta, s.thr, s.ll, s.f, s.fstar})
        if (p) dev_free(p);
    s = gpirt_sampler::DrawStore();
}
int64_t gpirt_store_sample_bytes(gpirt_sampler* h, int store_f, int store_fstar) {
    if (!h) return 0;
    int64_t b = 8LL * (h->n + 3LL * h->m_total + (int64_t)(h->C + 1) * h->m_total + 1);
    if (store_f) b += 8LL * h->n * h->m_total;
    if (store_fstar) b += 8LL * NGRID * h->m_total;
    return b;
}
int gpirt_store_open(gpirt_sampler* h, int capacity, int store_f, int store_fstar) {
    CHECK_H(h);
    if (h->world != 1 && (store_f || store_fstar))  // theta, beta, thresholds and the item terms are complete on every rank after the exchanges
        return fail(GPIRT_ERR_UNSUPPORTED, "draw store of f / f* on an item shard: not built (each rank holds its own items only)");
    if (h->cfg.p_beta != 3) return fail(GPIRT_ERR_SHAPE, "draw store needs 3 mean coefficients per item");
    if (capacity < 1) return fail(GPIRT_ERR_INVALID, "store_open: capacity %d", capacity);
    CU(cudaSetDevice(h->cfg.device));
    store_free(h);
    auto& s = h->store;
    const size_t cap = (size_t)capacity, m = (size_t)h->m_total;
    CU(dev_malloc((void**)&s.theta, cap * h->n * 8));
    CU(dev_malloc((void**)&s.beta, cap * 3 * m * 8));
    CU(dev_malloc((void**)&s.thr, cap * (h->C + 1) * m * 8));
    CU(dev_malloc((void**)&s.ll, cap * 8));
    if (store_f) CU(dev_malloc((void**)&s.f, cap * h->n * m * 8));
    if (store_fstar) CU(dev_malloc((void**)&s.fstar, cap * NGRID * m * 8));
    s.cap = capacity;
    s.count = 0;
    return GPIRT_OK;
}
int gpirt_store_push(gpirt_sampler* h) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    auto& s = h->store;
    if (s.cap < 1) return fail(GPIRT_ERR_INVALID, "store_push: no store open");
    if (s.count >= s.cap) return fail(GPIRT_ERR_INVALID, "store_push: store full (%d samples): flush first", s.cap);
    cudaStream_t st = h->stream;
    const size_t k = (size_t)s.count, m = (size_t)h->m_total;
    const int C1 = h->C + 1;
    // theta, beta, thresholds, f and f* are written by main-stream kernels only (the side stream owns L, L^-1, W, K_post)
    const int elems = std::max(h->n, std::max(3, C1) * h->m_total);
    // the total log-likelihood: the per-item terms of the sweep's last step (src/gpirtMCMC.cpp:324-331)
    k_store_sample<<<(elems + 255) / 256 + 1, 256, 0, st>>>(h->theta, h->n, h->beta, h->m_total, h->thr, C1, h->item_ll, s.theta + k * h->n,
                                                            s.beta + k * 3 * m, s.thr + k * C1 * m, s.ll + k);
    ++g_kernel_launches;
    CU(cudaGetLastError());
    if (s.f)
        CU(cudaMemcpy2DAsync(s.f + k * h->n * m, (size_t)h->n * 8, h->f, (size_t)h->ldn * 8, (size_t)h->n * 8, h->m_total, cudaMemcpyDeviceToDevice, st));
    if (s.fstar)  // the table is m x NGRID (ld_chunk) column-major in (item, grid point): back to NGRID x m
        CU(launch_transpose(st, h->fstar, h->ld_chunk, h->m_total, NGRID, s.fstar + k * NGRID * m, NGRID));
    s.count += 1;
    return GPIRT_OK;
}
int gpirt_store_count(gpirt_sampler* h, int* count, int* capacity) {
    CHECK_H(h);
    if (count) *count = h->store.count;
    if (capacity) *capacity = h->store.cap;
    return GPIRT_OK;
}
int gpirt_store_flush_slice(gpirt_sampler* h, int n_samples, int first, int horizon, int slice, double* theta_draws, double* beta_draws,
                            double* threshold_draws, double* ll_draws, double* f_draws, double* fstar_draws) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    auto& s = h->store;
    if (s.cap < 1) return fail(GPIRT_ERR_INVALID, "store_flush: no store open");
    if (first < 0 || first + s.count > n_samples) return fail(GPIRT_ERR_INVALID, "store_flush: samples %d..%d outside 0..%d", first, first + s.count, n_samples);
    if (horizon < 1 || slice < 0 || slice >= horizon) return fail(GPIRT_ERR_INVALID, "store_flush: slice %d of %d", slice, horizon);
    TRY(check_flags(h));  // one synchronisation for everything pushed since the last flush
    const size_t cnt = (size_t)s.count, m = (size_t)h->m_total, n = (size_t)h->n, C1 = (size_t)h->C + 1, H = (size_t)horizon, sl = (size_t)slice;
    if (cnt == 0) return GPIRT_OK;
    if (theta_draws) {  // cube n_samples x n x horizon: sample index is the fast one
        std::vector<double> t(cnt * n);
        CU(cudaMemcpy(t.data(), s.theta, t.size() * 8, cudaMemcpyDeviceToHost));
        double* base = theta_draws + sl * (size_t)n_samples * n;
        for (size_t i = 0; i < n; ++i)
            for (size_t k = 0; k < cnt; ++k) base[(size_t)first + k + i * (size_t)n_samples] = t[k * n + i];
    }
    // per sample a cube of `horizon` slices: slice `slice` of sample first + k sits at ((first + k) * horizon + slice) * slice_size
    auto rows = [&](double* dst, const double* src, size_t per) -> int {
        if (H == 1) { CU(cudaMemcpy(dst + (size_t)first * per, src, cnt * per * 8, cudaMemcpyDeviceToHost)); return GPIRT_OK; }
        CU(cudaMemcpy2D(dst + ((size_t)first * H + sl) * per, H * per * 8, src, per * 8, per * 8, cnt, cudaMemcpyDeviceToHost));
        return GPIRT_OK;
    };
    if (beta_draws) TRY(rows(beta_draws, s.beta, 3 * m));
    if (threshold_draws) TRY(rows(threshold_draws, s.thr, C1 * m));
    if (f_draws && s.f) TRY(rows(f_draws, s.f, n * m));
    if (fstar_draws && s.fstar) TRY(rows(fstar_draws, s.fstar, (size_t)NGRID * m));
    if (ll_draws) {  // the reference's ll is one sum over sessions and items (src/gpirtMCMC.cpp:324-331): slice 0 sets, later slices add
        std::vector<double> l(cnt);
        CU(cudaMemcpy(l.data(), s.ll, cnt * 8, cudaMemcpyDeviceToHost));
        for (size_t k = 0; k < cnt; ++k) ll_draws[(size_t)first + k] = (slice == 0 ? 0.0 : ll_draws[(size_t)first + k]) + l[k];
    }
    s.count = 0;
    return GPIRT_OK;
}
int gpirt_store_flush(gpirt_sampler* h, int n_samples, int first, double* theta_draws, double* beta_draws, double* threshold_draws,
                      double* ll_draws, double* f_draws, double* fstar_draws) {
    return gpirt_store_flush_slice(h, n_samples, first, 1, 0, theta_draws, beta_draws, threshold_draws, ll_draws, f_draws, fstar_draws);
}
int gpirt_store_close(gpirt_sampler* h) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaStreamSynchronize(h->stream));
    store_free(h);
    return GPIRT_OK;
}

int gpirt_get_L(gpirt_sampler* h, double* L) {
    CHECK_H(h);
    if (!L) return fail(GPIRT_ERR_INVALID, "get_L: NULL output");
    CU(cudaSetDevice(h->cfg.device));
    if (!h->chol_valid) return fail(GPIRT_ERR_INVALID, "get_L: the Cholesky cache is stale (theta changed); call gpirt_update_cholesky_cache");
    TRY(check_flags(h));
    CU(cudaMemcpy2DAsync(L, (size_t)h->n * 8, h->L, (size_t)h->ldn * 8, (size_t)h->n * 8, h->n, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return GPIRT_OK;
}

int gpirt_get_diag(gpirt_sampler* h, int block, int32_t* evals, double* margin) {
    CHECK_H(h);
    if (block < 0 || block > 4) return fail(GPIRT_ERR_INVALID, "get_diag: block %d", block);
    CU(cudaSetDevice(h->cfg.device));
    const size_t units = (block == GPIRT_BLOCK_THETA) ? h->n : h->m_total;
    if (evals) CU(cudaMemcpyAsync(evals, h->evals[block], units * 4, cudaMemcpyDeviceToHost, h->stream));
    if (margin) CU(cudaMemcpyAsync(margin, h->margin[block], units * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return GPIRT_OK;
}

}  // extern "C"

// ---- subsystem internals -------------------------------------------------------------------------
static ItemView item_view(gpirt_sampler* h) {
    ItemView v{};
    v.n = h->n; v.ldn = h->ldn; v.m_local = h->m_local; v.item0 = h->item0; v.m_total = h->m_total; v.C = h->C; v.p_beta = h->cfg.p_beta;
    v.y_im = h->y_im; v.ldy = h->ldy; v.theta = h->theta; v.beta = h->beta; v.thr = h->thr; v.f = h->f; v.nu = h->nu; v.sds = h->sds;
    v.z = nullptr; v.z_stride = 0; v.us.u = nullptr; v.us.stride = 0; v.key = h->key();
    v.item_ll = h->item_ll; v.flags = h->flags; v.screen = h->screen; v.early_exit = h->screen && h->early_exit; v.exact_evals = nullptr;
    v.mu_rows = h->mu_rows;
    return v;
}

// uploads the supplied streams of one block call; returns device pointers (or NULL for Philox)
static int stage_streams(gpirt_sampler* h, const gpirt_streams* s, size_t units, size_t nz, const double** dz, long long* zs,
                         UniformSrc* us) {
    *dz = nullptr; *zs = 0; us->u = nullptr; us->stride = 0;
    if (!s) return GPIRT_OK;
    if (s->z && nz > 0) {
        if (s->z_stride < (long long)nz) return fail(GPIRT_ERR_SHAPE, "z_stride %lld < %zu normals per unit", (long long)s->z_stride, nz);
        TRY(upload(h, &h->d_z, &h->d_z_cap, s->z, units * (size_t)s->z_stride));
        *dz = h->d_z; *zs = s->z_stride;
    }
    if (s->u) {
        if (s->u_stride < 2) return fail(GPIRT_ERR_SHAPE, "u_stride %lld < 2", (long long)s->u_stride);
        TRY(upload(h, &h->d_u, &h->d_u_cap, s->u, units * (size_t)s->u_stride));
        us->u = h->d_u; us->stride = s->u_stride;
    }
    return GPIRT_OK;
}

static int need_data(gpirt_sampler* h) {
    if (!h->have_data) return fail(GPIRT_ERR_INVALID, "no response data: call gpirt_set_data_* first");
    return GPIRT_OK;
}

// in-chain exchange helpers (side stream): every rank runs the same sequence of exchange points, so epochs advance in lockstep without a host handshake
static double* chain_peer_ptr(gpirt_sampler* h, int p, int which, const double* mine, const double* my_base) {
    return (double*)(h->chain_peer[p] + h->chain_off[p][which]) + (mine - my_base);
}
static int chain_signal(gpirt_sampler* h, cudaStream_t st, unsigned long long epoch) {  // publish `epoch` in this rank's word on every peer
    PeerPush pp{};
    for (int p = 0; p < h->world; ++p)
        if (p != h->rank) pp.flag[pp.count++] = (unsigned long long*)(h->chain_peer[p] + h->chain_off[p][gpirt_sampler::OFF_CFLAGS]) + h->rank;
    CU(launch_signal_peers(st, pp, epoch));
    return GPIRT_OK;
}
static int chain_wait(gpirt_sampler* h, cudaStream_t st, unsigned long long epoch, int only) {
    CU(launch_wait_peers(st, h->flags_c, h->world, h->rank, epoch, h->flags, only));
    return GPIRT_OK;
}
static int chain_push(gpirt_sampler* h, cudaStream_t st, int which, const double* my_base, const double* block, int ld, int rows, int cols) {
    PeerBlockPush pb{};
    for (int p = 0; p < h->world; ++p)
        if (p != h->rank) pb.dst[pb.count++] = chain_peer_ptr(h, p, which, block, my_base);
    CU(launch_push_block(st, block, ld, rows, cols, pb));
    return GPIRT_OK;
}


// The Cholesky's exchange points (linalg.cuh CholDist): block b of 256 columns belongs to rank b % world; its owner pushes the updated block to every peer.
struct ChainCholCtx { gpirt_sampler* h; unsigned long long base; };
static cudaError_t chain_chol_push(void* ctx, cudaStream_t st, int block_id, double* blk, int ld, int rows, int cols) {
    ChainCholCtx* c = (ChainCholCtx*)ctx;
    if (chain_push(c->h, st, gpirt_sampler::OFF_L, c->h->L, blk, ld, rows, cols) != GPIRT_OK) return cudaErrorUnknown;
    return chain_signal(c->h, st, c->base + (unsigned long long)block_id) == GPIRT_OK ? cudaSuccess : cudaErrorUnknown;
}
static cudaError_t chain_chol_wait(void* ctx, cudaStream_t st, int block_id, int owner) {
    ChainCholCtx* c = (ChainCholCtx*)ctx;
    return chain_wait(c->h, st, c->base + (unsigned long long)block_id, owner) == GPIRT_OK ? cudaSuccess : cudaErrorUnknown;
}

// trtri's exchange points (linalg.cuh TrtriDist)
static cudaError_t chain_trtri_push(void* ctx, cudaStream_t st, double* blk, int ld, int rows, int cols) {
    gpirt_sampler* h = (gpirt_sampler*)ctx;
    return chain_push(h, st, gpirt_sampler::OFF_LINV, h->Linv, blk, ld, rows, cols) == GPIRT_OK ? cudaSuccess : cudaErrorUnknown;
}
static cudaError_t chain_trtri_sync(void* ctx, cudaStream_t st) {
    gpirt_sampler* h = (gpirt_sampler*)ctx;
    const unsigned long long e = ++h->chain_epoch;
    return chain_signal(h, st, e) == GPIRT_OK && chain_wait(h, st, e, -1) == GPIRT_OK ? cudaSuccess : cudaErrorUnknown;
}

static int do_update_cholesky(gpirt_sampler* h, cudaStream_t st) {  // src/cholesky-cache.cpp:20-24
    SEC_ST(h, st, "K_build", CU(build_K(st, h->theta, h->n, h->theta, h->n, h->sds0, h->L, h->ldn, 1e-6)));
    // row milestones (nu = L Z in slices under the factorisation) only exist on the per-panel path: the persistent kernel finishes all rows at once, and
    // one launch of the product reads Z once instead of once per slice (ncu: 382 MB against 305 MB of DRAM traffic for the five slices)
    const bool marks = st == h->side && h->L_rows.count > 0 && !chol_is_persistent(h->n, h->plan_n.state ? &h->plan_n : nullptr);
    // Distributed chain (every rank is at this point of the same sweep: launch_theta_chain): first everything this rank writes into the matrices its peers
    // push into — raw K into L, k* into W — then a barrier, so that no pushed block can be overwritten by its receiver's own construction.
    ChainCholCtx cctx{h, 0};
    CholDist cd;
    const bool distributed = h->dist_chain && st == h->side && h->world > 1;
    if (distributed) {
        SEC_ST(h, st, "K_build", CU(build_K(st, h->theta, h->n, h->tstar, NGRID, h->sds0, h->W, h->ldn, 0.0)));
        h->kstar_built = true;
        const unsigned long long e = ++h->chain_epoch;
        TRY(chain_signal(h, st, e));
        TRY(chain_wait(h, st, e, -1));
        cctx.base = h->chain_epoch;
        h->chain_epoch += (unsigned long long)chol_dist_blocks(h->n);
        cd.rank = h->rank; cd.world = h->world; cd.ctx = &cctx; cd.push = chain_chol_push; cd.wait = chain_chol_wait;
    }
    SEC_ST(h, st, "chol_K", CU(chol_lower_blocked(st, h->L, h->ldn, h->n, &h->flags->chol_info, h->chol_ws, marks ? &h->L_rows : nullptr,
                                                      h->chol_aux.helper ? &h->chol_aux : nullptr, h->plan_n.state ? &h->plan_n : nullptr,
                                                      distributed ? &cd : nullptr)));
    h->pending_rows = marks;
    h->chol_valid = true;
    h->shared_valid = false;
    return GPIRT_OK;
}
static int do_update_cholesky(gpirt_sampler* h) {
    TRY(quiesce(h));
    return do_update_cholesky(h, h->stream);
}

static int do_draw_f(gpirt_sampler* h, const gpirt_streams* s) {
    if (!h->chol_valid) TRY(do_update_cholesky(h));  // the reference refactorises here (src/draw-f.cpp:59); same theta, same L
    // chol(K) still in flight on the side stream: nu = L Z is launched in row slices, each as soon as its rows of L are final (row block i of
    // L is complete after panel i), so most of the product runs UNDER the factorisation instead of after it. Same tiles, same k ranges:
    // the slices give the bits of the single launch.
    const bool sliced = h->pending_L && h->pending_rows && !(s && s->nu);
    if (!sliced) TRY(wait_L(h));
    cudaStream_t st = h->stream;
    ItemView v = item_view(h);
    const double* dz; long long zs;
    gpirt_streams s_u{};  // only the uniforms go through stage_streams; z / nu are matrices handled below
    if (s) { s_u.u = s->u; s_u.u_stride = s->u_stride; }
    TRY(stage_streams(h, &s_u, h->m_total, 0, &dz, &zs, &v.us));
    if (h->m_local > 0) {
        if (s && s->nu) {
            CU(cudaMemcpy2DAsync(h->nu, (size_t)h->ldn * 8, s->nu + (size_t)h->item0 * h->n, (size_t)h->n * 8, (size_t)h->n * 8, h->m_local,
                                 cudaMemcpyHostToDevice, st));
        } else {
            if (s && s->z) {
                if (s->z_stride < h->n) return fail(GPIRT_ERR_SHAPE, "draw_f: z_stride %lld < n", (long long)s->z_stride);
                CU(cudaMemcpy2DAsync(h->zbuf, (size_t)h->ldz * 8, s->z + (size_t)h->item0 * s->z_stride, (size_t)s->z_stride * 8,
                                     (size_t)h->n * 8, h->m_local, cudaMemcpyHostToDevice, st));
            } else {
                SEC(h, "gen_Z", CU(launch_gen_normals(st, h->zbuf, h->ldz, h->n, h->m_local, h->key(), ST_F_Z, h->item0)));
            }
            // nu = L * Z for every local item at once (rmvnorm_threadsafe per item, src/mvnormal.h:17-24)
            int r0 = 0;
            const int n_slices = sliced ? h->L_rows.count : 1;
            for (int c = 0; c < n_slices; ++c) {
                const int r1 = sliced ? h->L_rows.rows[c] : h->n;
                if (sliced) CU(cudaStreamWaitEvent(st, h->L_rows.ev[c], 0));
                GemmArgs g{};
                g.M = r1 - r0; g.N = h->m_local; g.K = h->n; g.m_off = r0;
                g.A = h->L + r0; g.lda = h->ldn; g.B = h->zbuf; g.ldb = h->ldz; g.C = h->nu + r0; g.ldc = h->ldn;
                g.alpha = 1.0; g.beta = 0.0; g.a_lower = 1;
                SEC(h, "gemm_LZ", CU(gemm_f64(st, false, false, g)));
                r0 = r1;
            }
        }
    }
    if (sliced) h->pending_rows = false;  // pending_L stays set: whoever needs ALL of chol's work (the zeroed upper triangle) still waits for ev_L
    v.evals = h->evals[GPIRT_BLOCK_F]; v.margin = h->margin[GPIRT_BLOCK_F]; v.exact_evals = h->exact[GPIRT_BLOCK_F];
    SEC(h, "ess_f", CU(launch_ess_items(st, IB_F, v)));
    return GPIRT_OK;
}

// The part of draw_fstar shared by all items (src/draw-fstar.cpp:26-36): k*, tmp = L^-1 k*, K_post = K** - tmp^T tmp + 1e-6 I,
// L_post = chol(K_post), and W = L^-T tmp = K^-1 k*. Depends on theta only, so a sweep runs it on the side stream.
static int prep_fstar_shared(gpirt_sampler* h, cudaStream_t st, cudaEvent_t ev_after_W = nullptr) {
    const int n = h->n, N = NGRID;
    const bool distributed = h->dist_chain && st == h->side && h->world > 1 && h->kstar_built && h->solver == 1;
    if (!h->kstar_built) SEC_ST(h, st, "K_build", CU(build_K(st, h->theta, n, h->tstar, N, h->sds0, h->W, h->ldn, 0.0)));
    h->kstar_built = false;
    const double* tmp = h->W;  // L^-1 k*
    bool forked_w = false;
    cudaEvent_t w_done = nullptr;
    if (h->solver == 1) {
        // L^-1 explicitly: every diagonal block inverted in parallel, log2(n/64) batched merge levels, then the two solves are GEMMs
        TrtriDist td;
        td.rank = h->rank; td.world = h->world; td.ctx = h; td.push = chain_trtri_push; td.sync = chain_trtri_sync;
        SEC_ST(h, st, "trtri", CU(trtri_lower(st, h->L, h->ldn, n, h->Linv, h->ldn, h->tri_scratch, distributed ? &td : nullptr)));
        // distributed: the grid points' columns are split over the ranks in runs of whole 64-column tiles (same tiles, same k ranges as the one launch);
        // every rank forms its columns of tmp = L^-1 k* and of W = L^-T tmp and pushes both to its peers
        int c_lo = 0, c_hi = N;
        if (distributed) kstar_dist_cols(N, h->world, h->rank, &c_lo, &c_hi);
        GemmArgs g{};
        g.M = n; g.N = c_hi - c_lo; g.K = n; g.A = h->Linv; g.lda = h->ldn; g.B = h->W + (size_t)c_lo * h->ldn; g.ldb = h->ldn;
        g.C = h->W2 + (size_t)c_lo * h->ldn; g.ldc = h->ldn;
        g.alpha = 1.0; g.beta = 0.0; g.a_lower = 1;
        SEC_ST(h, st, "gemm_tmp", CU(gemm_f64(st, false, false, g)));
        tmp = h->W2;
        // W = L^-T tmp = K^-1 k*: the per-item mean k*^T (L^-T L^-1 f_j) (:51-52) becomes one contraction F^T W. Formed BEFORE K_post and its
        // Cholesky: the consumer's first product (F^T W) then starts while the latency-bound chol(K_post) is still running.
        GemmArgs w{};  // W = (L^-1)^T * tmp; op(A) = Linv^T is upper triangular
        w.M = n; w.N = c_hi - c_lo; w.K = n; w.A = h->Linv; w.lda = h->ldn; w.B = h->W2 + (size_t)c_lo * h->ldn; w.ldb = h->ldn;
        w.C = h->W + (size_t)c_lo * h->ldn; w.ldc = h->ldn;
        w.alpha = 1.0; w.beta = 0.0; w.a_upper = 1;
        // W and K_post both need tmp only: inside a sweep W goes to a second side stream and K_post with its latency-bound Cholesky starts at once,
        // instead of after W (the f* noise term was waiting 0.2 ms for chol(K_post) at C3, tools/timeline_c3.py)
        static const int fork_env = getenv("GPIRT_FORK_W") ? atoi(getenv("GPIRT_FORK_W")) : 1;
        const bool fork_w = fork_env && !distributed && st == h->side && h->side2 && ev_after_W;
        cudaStream_t sw = fork_w ? h->side2 : st;
        if (fork_w) {
            CU(cudaEventRecord(h->ev_tmp, st));
            CU(cudaStreamWaitEvent(sw, h->ev_tmp, 0));
        }
        SEC_ST(h, sw, "gemm_W", CU(gemm_f64(sw, true, false, w)));
        if (fork_w) {
            CU(cudaEventRecord(ev_after_W, sw));
            w_done = ev_after_W;
            ev_after_W = nullptr;  // recorded; `st` joins it at the end, so that whatever follows on `st` (ev_shared) covers W as before
            forked_w = true;
        }
        if (distributed) {
            if (c_hi > c_lo) {
                TRY(chain_push(h, st, gpirt_sampler::OFF_W2, h->W2, h->W2 + (size_t)c_lo * h->ldn, h->ldn, n, c_hi - c_lo));
                TRY(chain_push(h, st, gpirt_sampler::OFF_W, h->W, h->W + (size_t)c_lo * h->ldn, h->ldn, n, c_hi - c_lo));
            }
            const unsigned long long e = ++h->chain_epoch;
            TRY(chain_signal(h, st, e));
            TRY(chain_wait(h, st, e, -1));
        }
        if (ev_after_W) CU(cudaEventRecord(ev_after_W, st));
    } else {
        SEC_ST(h, st, "trsm_kstar", CU(trsm_left_lower(st, h->L, h->ldn, n, h->W, h->ldn, N)));
    }
    CU(build_K(st, h->tstar, N, h->tstar, N, h->sds0, h->Kpost, h->ldN, 0.0));
    {
        GemmArgs g{};
        g.M = N; g.N = N; g.K = n; g.A = tmp; g.lda = h->ldn; g.B = tmp; g.ldb = h->ldn; g.C = h->Kpost; g.ldc = h->ldN;
        g.alpha = -1.0; g.beta = 1.0; g.lower_only = 1;
        SEC_ST(h, st, "gemm_Kpost", CU(gemm_f64(st, true, false, g)));
    }
    k_add_diag<<<(N + 255) / 256, 256, 0, st>>>(h->Kpost, h->ldN, N, 1e-6);
    ++g_kernel_launches;
    CU(cudaGetLastError());
    const bool post_marks = st == h->side && h->Lpost_rows.count > 0 && !chol_is_persistent(N, h->plan_N.state ? &h->plan_N : nullptr);
    SEC_ST(h, st, "chol_Kpost", CU(chol_lower_blocked(st, h->Kpost, h->ldN, N, &h->flags->chol_post_info, h->chol_ws, post_marks ? &h->Lpost_rows : nullptr,
                                                          nullptr, h->plan_N.state ? &h->plan_N : nullptr)));
    h->pending_post_rows = post_marks;
    if (forked_w) CU(cudaStreamWaitEvent(st, w_done, 0));
    if (h->solver != 1) {  // substitution overwrites tmp in place, so it comes after K_post
        SEC_ST(h, st, "trsm_W", CU(trsm_left_lower_t(st, h->L, h->ldn, n, h->W, h->ldn, N)));
        if (ev_after_W) CU(cudaEventRecord(ev_after_W, st));
    }
    h->shared_valid = true;
    return GPIRT_OK;
}

// After a new theta is committed inside a sweep: K, chol(K) and the shared draw_fstar terms of the NEXT sweep start on the side
// stream right away (the reference refactorises at this point too, src/gpirtMCMC.cpp:314). The main stream joins at first use.
static int launch_theta_chain(gpirt_sampler* h) {
    // serial order when profiling (update_cholesky runs where the reference has it) unless the timeline of the overlapped
    // schedule itself is wanted (profiling level 2: section times then include waiting and sharing the SMs)
    if (!h->overlap || (h->profiling && h->profile_level < 2)) return GPIRT_OK;
    CU(cudaEventRecord(h->ev_fork, h->stream));
    CU(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
    TRY(do_update_cholesky(h, h->side));
    CU(cudaEventRecord(h->ev_L, h->side));
    h->pending_L = true;
    // (Launching this second half only behind the next sweep's L Z product, so that its tensor work runs under the ESS kernel instead of sharing the DMMA
    // pipe with the product, was measured: 4.88 against 4.61 ms per sweep at C3 — the chain then ends after the f* draw needs it.)
    TRY(prep_fstar_shared(h, h->side, h->ev_W));
    h->pending_W = true;
    CU(cudaEventRecord(h->ev_shared, h->side));
    h->pending_shared = true;
    return GPIRT_OK;
}

static int do_draw_fstar(gpirt_sampler* h, const gpirt_streams* s, bool push_to_peers = false) {  // src/draw-fstar.cpp:18-57
    if (!h->chol_valid) TRY(do_update_cholesky(h));
    if (!h->shared_valid) {
        TRY(quiesce(h));
        TRY(prep_fstar_shared(h, h->stream));
    }
    TRY(wait_W(h));  // W only; L_post is waited for where it is used
    cudaStream_t st = h->stream;
    const int n = h->n, N = NGRID;
    if (h->m_local == 0) return quiesce(h);
    double* chunk = h->fstar + (size_t)h->rank * NGRID * h->ld_chunk;
    {
        GemmArgs g{};
        g.M = h->m_local; g.N = N; g.K = n; g.A = h->f; g.lda = h->ldn; g.B = h->W; g.ldb = h->ldn; g.C = chunk; g.ldc = h->ld_chunk;
        g.alpha = 1.0; g.beta = 0.0;
        SEC(h, "gemm_FtW", CU(gemm_f64(st, true, false, g)));
    }
    if (s && s->z) {
        if (s->z_stride < N) return fail(GPIRT_ERR_SHAPE, "draw_fstar: z_stride %lld < %d", (long long)s->z_stride, N);
        CU(cudaMemcpy2DAsync(h->zbuf, (size_t)h->ldz * 8, s->z + (size_t)h->item0 * s->z_stride, (size_t)s->z_stride * 8, (size_t)N * 8,
                             h->m_local, cudaMemcpyHostToDevice, st));
    } else {
        SEC(h, "gen_Z", CU(launch_gen_normals(st, h->zbuf, h->ldz, N, h->m_local, h->key(), ST_FSTAR_Z, h->item0)));
    }
    // + (L_post Z*)^T : op(A) = Z*^T, op(B) = L_post^T (upper triangular). Grid-point block c of the result needs rows [0, end of block c) of
    // L_post only, which are final after panel c of chol(K_post): while that factorisation is still running on the side stream the product
    // goes out in column slices behind its milestones (same tiles, same k ranges, same bits).
    const bool post_sliced = h->pending_shared && h->pending_post_rows;
    if (!post_sliced) TRY(wait_shared(h));
    {
        int c0 = 0;
        const int n_slices = post_sliced ? h->Lpost_rows.count : 1;
        for (int c = 0; c < n_slices; ++c) {
            const int c1 = post_sliced ? h->Lpost_rows.rows[c] : N;
            if (post_sliced) CU(cudaStreamWaitEvent(st, h->Lpost_rows.ev[c], 0));
            GemmArgs g{};
            g.M = h->m_local; g.N = c1 - c0; g.K = N; g.n_off = c0;
            g.A = h->zbuf; g.lda = h->ldz; g.B = h->Kpost + c0; g.ldb = h->ldN; g.C = chunk + (size_t)c0 * h->ld_chunk; g.ldc = h->ld_chunk;
            g.alpha = 1.0; g.beta = 1.0; g.b_upper = 1;
            SEC(h, "gemm_ZLpost", CU(gemm_f64(st, true, true, g)));
            c0 = c1;
        }
    }
    if (post_sliced) h->pending_post_rows = false;  // pending_shared stays set (the tail of chol(K_post): zeroing the upper triangle)
    if (push_to_peers && h->world > 1 && h->peers_attached == h->world - 1) {
        // the exchange that feeds the theta step, fused into the last kernel of the f* draw: finished values go straight into every
        // peer's table over NVLink, then the epoch is published in the peers' flag words (gpirt_peer_wait on the other side)
        PeerPush pp{};
        for (int p = 0; p < h->world; ++p) {  // a peer's arena is laid out for ITS shard sizes: its own offsets, chunk / word `rank` within them
            if (p == h->rank) continue;
            pp.chunk[pp.count] = (double*)(h->peer_arena[p] + h->peer_fstar_off[p]) + (size_t)h->rank * NGRID * h->ld_chunk;
            pp.flag[pp.count] = (unsigned long long*)(h->peer_arena[p] + h->peer_flags_off[p]) + h->rank;
            ++pp.count;
        }
        h->push_epoch += 1;
        SEC(h, "add_mustar_push", CU(launch_add_mustar_push(st, chunk, h->ld_chunk, h->m_local, h->item0, h->beta, h->cfg.p_beta, pp)));
        CU(launch_signal_peers(st, pp, h->push_epoch));
        return GPIRT_OK;
    }
    SEC(h, "add_mustar", CU(launch_add_mustar(st, chunk, h->ld_chunk, h->m_local, h->item0, h->beta, h->cfg.p_beta)));
    return GPIRT_OK;
}

static int need_theta_tab(gpirt_sampler* h, size_t bytes) {  // the table of the respondent-sharded and of the item-sharded theta step share one block
    if (h->theta_tab && h->theta_tab_bytes >= bytes) return GPIRT_OK;
    if (h->theta_tab) {
        CU(cudaStreamSynchronize(h->stream));
        CU(dev_free(h->theta_tab));
        h->theta_tab = nullptr;
    }
    CU(dev_malloc((void**)&h->theta_tab, bytes));
    h->theta_tab_bytes = bytes;
    return GPIRT_OK;
}

// table_only != nullptr: only prepare the per-sweep term table (or decide against it) and report it — the joint step of a panel launches its own kernel
static int do_draw_theta(gpirt_sampler* h, const gpirt_streams* s, const double** table_only = nullptr) {
    ThetaView v{};
    v.n = h->n; v.resp0 = h->resp0; v.n_local = h->n_local; v.m_total = h->m_total; v.m_chunk = h->m_chunk; v.ld_chunk = h->ld_chunk;
    v.world = h->world; v.C = h->C; v.y_rm = h->y_rm; v.ldm = h->ldm; v.theta = h->theta; v.theta_new = h->theta_x;
    v.fstar_all = h->fstar; v.chunk_elems = (long long)NGRID * h->ld_chunk; v.beta = h->beta; v.thr = h->thr_theta ? h->thr_theta : h->thr;
    v.theta_prior_sds = h->tps;
    v.regime = h->theta_regime; v.theta_os = h->cfg.theta_os; v.key = h->key();
    v.evals = h->evals[GPIRT_BLOCK_THETA]; v.margin = h->margin[GPIRT_BLOCK_THETA]; v.flags = h->flags;
    TRY(stage_streams(h, s, h->n, 1, &v.z, &v.z_stride, &v.us));
    // Every term the step can ask for is a function of (grid point, item, category) only: unless this rank has very few respondents, build
    // the table of all NGRID x C terms per item once per sweep instead of evaluating ~12 terms per (respondent, item) on the fly. Measured
    // on B200 (profiles/r01_scaling.txt): the table kernel makes ~190 G terms/s (coalesced, every SM busy), the on-the-fly path ~21 G terms/s
    // (one CTA per respondent, gathers across the f* chunks, two Phi and a log per term) -> a term costs ~9x more on the fly.
    const double table_cost = (double)NGRID * h->C, direct_cost = (double)h->n_local * 12.0 * 9.0;
    const size_t table_bytes = (size_t)NGRID * h->m_total * h->C * 8;
    v.table = nullptr;
    v.table32 = nullptr;
    v.exact_evals = h->exact[GPIRT_BLOCK_THETA];
    if (direct_cost > table_cost && table_bytes <= ((size_t)12 << 30)) {
        TRY(need_theta_tab(h, table_bytes));
        // The bounds add up over the items (~3e-6 per term): at m = 5000 0.8 % of the proposals stay undecided, at m = 40 000 (8 item shards of C3)
        // 4 %, and an undecided proposal costs two fp64 evaluations of all m items on ONE SM (75 us at m = 40 000) — measured at 8 GPUs the
        // screened step is then slower than the fp64 table (0.85 + 0.92 ms against 1.03 + 0.45 ms), so it is used up to 12 000 items.
        if (h->screen && !table_only && h->m_total <= 12000) {
            // fp32 screening (gpirt_set_screen): the table holds (fp32 term, error bound) pairs — the same bytes, a quarter of the
            // arithmetic to build — and the few proposals it cannot decide are evaluated in fp64 on the fly
            SEC(h, "theta_table", CU(launch_theta_table32(h->stream, v, (float2*)h->theta_tab)));
            v.table32 = (const float2*)h->theta_tab;
        } else {
            SEC(h, "theta_table", CU(launch_theta_table(h->stream, v, h->theta_tab)));
            v.table = h->theta_tab;
        }
    }
    if (table_only) { *table_only = v.table; return GPIRT_OK; }
    SEC(h, "ess_theta", CU(launch_ess_theta(h->stream, v)));
    return GPIRT_OK;
}

static int commit_theta(gpirt_sampler* h) {  // theta.swap(theta_new), src/gpirtMCMC.cpp:284
    TRY(quiesce(h));  // nothing of the previous theta's chain may still be reading it
    CU(cudaMemcpyAsync(h->theta, h->theta_x, (size_t)h->n * 8, cudaMemcpyDeviceToDevice, h->stream));
    h->chol_valid = h->shared_valid = false;
    return GPIRT_OK;
}

static int do_gather_f(gpirt_sampler* h) {
    if (h->m_local == 0) return GPIRT_OK;
    SEC(h, "gather_f", CU(launch_gather_f(h->stream, h->theta, h->n, h->fstar + (size_t)h->rank * NGRID * h->ld_chunk, h->ld_chunk, h->m_local, h->f, h->ldn)));
    return GPIRT_OK;
}

static int do_small_ess(gpirt_sampler* h, int which, const gpirt_streams* s) {
    ItemView v = item_view(h);
    const int blk = which == IB_BETA ? GPIRT_BLOCK_BETA : GPIRT_BLOCK_THR;
    const size_t nz = which == IB_BETA ? 3 : (size_t)(h->C - 1);
    TRY(stage_streams(h, s, h->m_total, nz, &v.z, &v.z_stride, &v.us));
    v.evals = h->evals[blk]; v.margin = h->margin[blk]; v.exact_evals = h->exact[blk];
    SEC(h, which == IB_BETA ? "ess_beta" : "ess_thr", CU(launch_ess_items(h->stream, which, v)));
    return GPIRT_OK;
}

static int do_item_ll(gpirt_sampler* h) {
    ItemView v = item_view(h);
    SEC(h, "item_ll", CU(launch_item_ll(h->stream, v)));
    return GPIRT_OK;
}

extern "C" {

#define ENTER(h)                         \
    CHECK_H(h);                          \
    CU(cudaSetDevice((h)->cfg.device));  \
    TRY(need_data(h))

int gpirt_update_cholesky_cache(gpirt_sampler* h) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    return do_update_cholesky(h);
}
int gpirt_draw_f(gpirt_sampler* h, const gpirt_streams* s) { ENTER(h); return do_draw_f(h, s); }
int gpirt_draw_fstar(gpirt_sampler* h, const gpirt_streams* s) { ENTER(h); return do_draw_fstar(h, s); }
int gpirt_draw_theta(gpirt_sampler* h, const gpirt_streams* s) {
    ENTER(h);
    TRY(do_draw_theta(h, s));
    if (h->world == 1) return commit_theta(h);  // world > 1: the caller all-gathers XCHG_THETA, then gpirt_sweep_phase_close commits
    return GPIRT_OK;
}
int gpirt_gather_f(gpirt_sampler* h) { ENTER(h); return do_gather_f(h); }
int gpirt_draw_beta(gpirt_sampler* h, const gpirt_streams* s) {
    ENTER(h);
    if (h->cfg.p_beta != 3) return fail(GPIRT_ERR_SHAPE, "draw_beta needs 3 mean coefficients per item (X is n x 3, src/gpirtMCMC.cpp:138-141)");
    return do_small_ess(h, IB_BETA, s);
}
int gpirt_draw_threshold(gpirt_sampler* h, const gpirt_streams* s) { ENTER(h); return do_small_ess(h, IB_THR, s); }

int gpirt_total_ll(gpirt_sampler* h, double* total) {
    CHECK_H(h);
    if (!total) return fail(GPIRT_ERR_INVALID, "total_ll: NULL output");
    CU(cudaSetDevice(h->cfg.device));
    CU(launch_ordered_sum(h->stream, h->item_ll, h->m_total, h->total_ll));
    CU(cudaMemcpyAsync(total, h->total_ll, 8, cudaMemcpyDeviceToHost, h->stream));
    return check_flags(h);
}

int gpirt_ll_bar(gpirt_sampler* h, double* per_item, double* total) {
    ENTER(h);
    TRY(do_item_ll(h));
    if (per_item) CU(cudaMemcpyAsync(per_item, h->item_ll, (size_t)h->m_total * 8, cudaMemcpyDeviceToHost, h->stream));
    if (total) return gpirt_total_ll(h, total);  // world > 1: only meaningful after the ITEM_LL exchange
    return check_flags(h);
}

// ---- sweep (src/gpirtMCMC.cpp:267-331) ----
int gpirt_sweep_phase_items(gpirt_sampler* h) {
    ENTER(h);
    TRY(do_draw_f(h, nullptr));
    return do_draw_fstar(h, nullptr, true);
}

// ---- f* exchange through peer memory (one process per GPU: cudaIpc mappings of the peers' arenas) ----
int gpirt_ipc_export(gpirt_sampler* h, void* handle64, int64_t* fstar_offset, int64_t* flags_offset) {
    CHECK_H(h);
    if (!handle64 || !fstar_offset || !flags_offset) return fail(GPIRT_ERR_INVALID, "ipc_export: NULL output");
    *fstar_offset = (int64_t)((char*)h->fstar - h->arena);
    *flags_offset = (int64_t)((char*)h->flags_x - h->arena);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaStreamSynchronize(h->stream));  // the arena's zero fill (flag words included) is complete before anyone maps it
    cudaIpcMemHandle_t mh;
    CU(cudaIpcGetMemHandle(&mh, h->arena));
    std::memcpy(handle64, &mh, 64);
    return GPIRT_OK;
}
int gpirt_ipc_attach(gpirt_sampler* h, int peer_rank, const void* handle64, int64_t fstar_offset, int64_t flags_offset) {
    CHECK_H(h);
    if (!handle64 || peer_rank < 0 || peer_rank >= h->world || peer_rank == h->rank || h->world > gpirt::MAX_PEERS + 1)
        return fail(GPIRT_ERR_INVALID, "ipc_attach: peer rank %d of a world of %d (at most %d ranks)", peer_rank, h->world, gpirt::MAX_PEERS + 1);
    if (h->peer_arena[peer_rank]) return fail(GPIRT_ERR_INVALID, "ipc_attach: rank %d is attached already", peer_rank);
    CU(cudaSetDevice(h->cfg.device));
    cudaIpcMemHandle_t mh;
    std::memcpy(&mh, handle64, 64);
    void* base = nullptr;
    CU(cudaIpcOpenMemHandle(&base, mh, cudaIpcMemLazyEnablePeerAccess));
    h->peer_arena[peer_rank] = (char*)base;
    h->peer_fstar_off[peer_rank] = (size_t)fstar_offset;
    h->peer_flags_off[peer_rank] = (size_t)flags_offset;
    h->peers_attached += 1;
    return GPIRT_OK;
}
int gpirt_ipc_detach(gpirt_sampler* h) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    TRY(quiesce(h));
    CU(cudaStreamSynchronize(h->stream));
    for (char*& pa : h->peer_arena)
        if (pa) { CU(cudaIpcCloseMemHandle(pa)); pa = nullptr; }
    h->peers_attached = 0;
    return GPIRT_OK;
}
int gpirt_peer_wait(gpirt_sampler* h) {
    ENTER(h);
    if (h->world == 1) return GPIRT_OK;
    if (h->peers_attached != h->world - 1) return fail(GPIRT_ERR_INVALID, "peer_wait: %d of %d peers attached", h->peers_attached, h->world - 1);
    CU(launch_wait_peers(h->stream, h->flags_x, h->world, h->rank, h->push_epoch, h->flags));
    return GPIRT_OK;
}
// ---- distributed theta-only chain: peers' chain buffers (one process per GPU: cudaIpc; one process, several GPUs: the pointers themselves) ----
int gpirt_chain_export(gpirt_sampler* h, void* handle64, int64_t* offsets6) {
    CHECK_H(h);
    if (!handle64 || !offsets6) return fail(GPIRT_ERR_INVALID, "chain_export: NULL output");
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaStreamSynchronize(h->stream));  // the arena's zero fill (flag words included) is complete before anyone maps it
    offsets6[gpirt_sampler::OFF_L] = (int64_t)((char*)h->L - h->arena);
    offsets6[gpirt_sampler::OFF_LINV] = (int64_t)((char*)h->Linv - h->arena);
    offsets6[gpirt_sampler::OFF_W] = (int64_t)((char*)h->W - h->arena);
    offsets6[gpirt_sampler::OFF_W2] = (int64_t)((char*)h->W2 - h->arena);
    offsets6[gpirt_sampler::OFF_CFLAGS] = (int64_t)((char*)h->flags_c - h->arena);
    offsets6[5] = (int64_t)(intptr_t)h->arena;  // meaningful inside the exporting process only
    cudaIpcMemHandle_t mh;
    CU(cudaIpcGetMemHandle(&mh, h->arena));
    std::memcpy(handle64, &mh, 64);
    return GPIRT_OK;
}
// handle64 != NULL: the peer lives in another process (IPC mapping); NULL: in this one (offsets6[5] is its arena; peer access between two GPUs is enabled
// here). Once world - 1 peers are attached and n >= 4096 the theta-only chain of a sweep is distributed (capi.cu).
int gpirt_chain_attach(gpirt_sampler* h, int peer_rank, const void* handle64, const int64_t* offsets6) {
    CHECK_H(h);
    if (!offsets6 || peer_rank < 0 || peer_rank >= h->world || peer_rank == h->rank || h->world > gpirt::MAX_PEERS + 1)
        return fail(GPIRT_ERR_INVALID, "chain_attach: peer rank %d of a world of %d (at most %d ranks)", peer_rank, h->world, gpirt::MAX_PEERS + 1);
    if (h->chain_peer[peer_rank]) return fail(GPIRT_ERR_INVALID, "chain_attach: rank %d is attached already", peer_rank);
    CU(cudaSetDevice(h->cfg.device));
    if (handle64) {
        if (h->peer_arena[peer_rank]) h->chain_peer[peer_rank] = h->peer_arena[peer_rank];  // mapped already for the f* push
        else {
            cudaIpcMemHandle_t mh;
            std::memcpy(&mh, handle64, 64);
            void* base = nullptr;
            CU(cudaIpcOpenMemHandle(&base, mh, cudaIpcMemLazyEnablePeerAccess));
            h->chain_peer[peer_rank] = (char*)base;
            h->chain_peer_ipc[peer_rank] = true;
        }
    } else {
        h->chain_peer[peer_rank] = (char*)(intptr_t)offsets6[5];
        cudaPointerAttributes pa{};
        CU(cudaPointerGetAttributes(&pa, h->chain_peer[peer_rank]));
        if (pa.type != cudaMemoryTypeDevice) return fail(GPIRT_ERR_INVALID, "chain_attach: offsets6[5] is not a device address of this process");
        if (pa.device != h->cfg.device) {  // this GPU's kernels store into the peer's arena over NVLink
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, h->cfg.device, pa.device));
            if (!can) return fail(GPIRT_ERR_INVALID, "chain_attach: device %d has no peer access to device %d", h->cfg.device, pa.device);
            cudaError_t pe = cudaDeviceEnablePeerAccess(pa.device, 0);
            if (pe == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
            else CU(pe);
        }
    }
    for (int k = 0; k < gpirt_sampler::N_CHAIN_OFFS; ++k) h->chain_off[peer_rank][k] = (size_t)offsets6[k];
    h->chain_attached += 1;
    static const int enabled = getenv("GPIRT_DIST_CHAIN") ? atoi(getenv("GPIRT_DIST_CHAIN")) : 1;
    static const int min_n = getenv("GPIRT_DIST_CHAIN_MIN_N") ? atoi(getenv("GPIRT_DIST_CHAIN_MIN_N")) : 4096;
    h->dist_chain = enabled && h->chain_attached == h->world - 1 && h->n >= min_n;
    return GPIRT_OK;
}
// Test hook (no device needed): rank `rank`'s share of the distributed theta-only chain for n respondents on `world` ranks — out[0..1] its columns of the
// grid points (L^-1 k*, K^-1 k*), out[2] the number of 256-column Cholesky blocks it owns, out[3] the number of split inversion levels, then for each
// split level (block widths ascending) its block width and its column range: out[4 + 3 l .. 6 + 3 l]. Returns the number of ints written.
int gpirt_debug_dist_plan(int n, int world, int rank, int32_t* out, int cap) {
    if (!out || n < 1 || world < 1 || rank < 0 || rank >= world || cap < 4) return fail(GPIRT_ERR_INVALID, "debug_dist_plan: bad arguments");
    int lo = 0, hi = 0;
    kstar_dist_cols((int)NGRID, world, rank, &lo, &hi);
    out[0] = lo; out[1] = hi;
    int owned = 0;
    for (int b = 0; b < chol_dist_blocks(n); ++b) owned += chol_dist_owner(b, world) == rank;
    out[2] = owned;
    int k = 4, levels = 0;
    for (int b = CHOL_NB; b < n; b *= 2) {
        if (!trtri_dist_level(b, world)) continue;
        if (k + 3 > cap) return fail(GPIRT_ERR_INVALID, "debug_dist_plan: cap too small");
        out[k++] = b; out[k++] = trtri_dist_cut(b, world, rank); out[k++] = trtri_dist_cut(b, world, rank + 1);
        ++levels;
    }
    out[3] = levels;
    return k;
}
int gpirt_chain_detach(gpirt_sampler* h) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    TRY(quiesce(h));
    CU(cudaStreamSynchronize(h->stream));
    for (int p = 0; p <= gpirt::MAX_PEERS; ++p) {
        if (h->chain_peer[p] && h->chain_peer_ipc[p] && h->chain_peer[p] != h->peer_arena[p]) CU(cudaIpcCloseMemHandle(h->chain_peer[p]));
        h->chain_peer[p] = nullptr;
        h->chain_peer_ipc[p] = false;
    }
    h->chain_attached = 0;
    h->dist_chain = false;
    return GPIRT_OK;
}

int gpirt_sweep_phase_theta(gpirt_sampler* h) { ENTER(h); return do_draw_theta(h, nullptr); }

// ---- item-sharded theta step (samplers.cuh ThetaItemsView): begin, then rounds of { partials, exchange THETA_PART, decide } until nobody is active ----
static ThetaItemsView theta_items_view(gpirt_sampler* h) {
    ThetaItemsView v{};
    v.n = h->n; v.m_local = h->m_local; v.C = h->C; v.world = h->world;
    v.y_rl = h->y_rl; v.ldml = h->ldml; v.table = h->theta_tab; v.theta = h->theta; v.theta_new = h->theta_x;
    v.theta_prior_sds = h->tps; v.regime = h->theta_regime; v.theta_os = h->cfg.theta_os;
    v.z = h->theta_z_dev; v.z_stride = h->theta_z_stride; v.us = h->theta_us; v.key = h->key();
    v.part_mine = h->theta_part + (size_t)h->rank * h->n * (THETA_BATCH + 1); v.part_all = h->theta_part;
    v.k_next = h->theta_knext; v.ll0 = h->theta_ll0; v.n_active = h->theta_nact;
    v.evals = h->evals[GPIRT_BLOCK_THETA]; v.margin = h->margin[GPIRT_BLOCK_THETA]; v.flags = h->flags;
    return v;
}
int gpirt_theta_items_begin(gpirt_sampler* h, const gpirt_streams* s) {
    ENTER(h);
    if (h->world < 2 || !h->y_rl) return fail(GPIRT_ERR_INVALID, "theta_items: the item-sharded theta step belongs to an item shard (world > 1)");
    TRY(stage_streams(h, s, h->n, 1, &h->theta_z_dev, &h->theta_z_stride, &h->theta_us));
    // the term table of the LOCAL items only: f* chunk, beta and thresholds of this rank (every other rank builds its own)
    const size_t table_bytes = (size_t)NGRID * std::max(h->m_local, 1) * h->C * 8;
    TRY(need_theta_tab(h, std::max(table_bytes, (size_t)NGRID * h->m_chunk * h->C * 8)));
    if (h->m_local > 0) {
        ThetaView t{};
        t.m_total = h->m_local; t.m_chunk = h->m_local; t.ld_chunk = h->ld_chunk; t.world = 1; t.C = h->C;
        t.fstar_all = h->fstar + (size_t)h->rank * NGRID * h->ld_chunk; t.chunk_elems = (long long)NGRID * h->ld_chunk;
        t.beta = h->beta + 3 * (size_t)h->item0;
        t.thr = (h->thr_theta ? h->thr_theta : h->thr) + (size_t)(h->C + 1) * h->item0;
        SEC(h, "theta_table", CU(launch_theta_table(h->stream, t, h->theta_tab)));
    }
    CU(cudaMemsetAsync(h->theta_knext, 0, (size_t)h->n * 4, h->stream));
    return GPIRT_OK;
}
int gpirt_theta_items_partials(gpirt_sampler* h) {
    ENTER(h);
    if (h->world < 2 || !h->theta_tab) return fail(GPIRT_ERR_INVALID, "theta_items_partials: call gpirt_theta_items_begin first");
    ThetaItemsView v = theta_items_view(h);
    SEC(h, "ess_theta", CU(launch_theta_partials(h->stream, v)));
    return GPIRT_OK;
}
int gpirt_theta_items_decide(gpirt_sampler* h, int32_t* n_active) {
    ENTER(h);
    if (h->world < 2 || !n_active) return fail(GPIRT_ERR_INVALID, "theta_items_decide: bad arguments");
    ThetaItemsView v = theta_items_view(h);
    CU(cudaMemsetAsync(h->theta_nact, 0, 4, h->stream));
    SEC(h, "theta_decide", CU(launch_theta_decide(h->stream, v)));
    CU(cudaMemcpyAsync(n_active, h->theta_nact, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));  // the caller's loop ends on this count
    return GPIRT_OK;
}
int gpirt_items_pack(gpirt_sampler* h, int unpack) {
    ENTER(h);
    if (h->world < 2) return GPIRT_OK;
    CU(launch_pack_items(h->stream, h->beta, h->thr, h->item_ll, h->C + 1, h->m_chunk, h->rank, h->world, h->items_packed, unpack));
    return GPIRT_OK;
}
int gpirt_sweep_phase_close(gpirt_sampler* h) {
    ENTER(h);
    TRY(commit_theta(h));
    TRY(launch_theta_chain(h));
    TRY(do_gather_f(h));
    TRY(do_small_ess(h, IB_BETA, nullptr));
    if (!h->chol_valid) TRY(do_update_cholesky(h));
    TRY(do_small_ess(h, IB_THR, nullptr));
    TRY(do_item_ll(h));
    h->iter += 1;
    return GPIRT_OK;
}
int gpirt_sweep(gpirt_sampler* h, int n_sweeps) {
    ENTER(h);
    if (h->world != 1) return fail(GPIRT_ERR_INVALID, "gpirt_sweep on an item shard: drive the three phases and the exchanges instead");
    if (h->cfg.p_beta != 3) return fail(GPIRT_ERR_SHAPE, "a sweep needs 3 mean coefficients per item");
    for (int k = 0; k < n_sweeps; ++k) {
        TRY(gpirt_sweep_phase_items(h));
        TRY(gpirt_sweep_phase_theta(h));
        TRY(gpirt_sweep_phase_close(h));
    }
    return GPIRT_OK;
}

int gpirt_exchange_buffer(gpirt_sampler* h, int which, void** dev_ptr, int64_t* chunk_bytes) {
    CHECK_H(h);
    if (!dev_ptr || !chunk_bytes) return fail(GPIRT_ERR_INVALID, "exchange_buffer: NULL output");
    switch (which) {
        case GPIRT_XCHG_FSTAR: *dev_ptr = h->fstar; *chunk_bytes = (int64_t)NGRID * h->ld_chunk * 8; break;
        case GPIRT_XCHG_BETA: *dev_ptr = h->beta; *chunk_bytes = (int64_t)3 * h->m_chunk * 8; break;
        case GPIRT_XCHG_THR: *dev_ptr = h->thr; *chunk_bytes = (int64_t)(h->C + 1) * h->m_chunk * 8; break;
        case GPIRT_XCHG_THETA: *dev_ptr = h->theta_x; *chunk_bytes = (int64_t)h->n_chunk * 8; break;
        case GPIRT_XCHG_ITEM_LL: *dev_ptr = h->item_ll; *chunk_bytes = (int64_t)h->m_chunk * 8; break;
        case GPIRT_XCHG_THETA_PART: *dev_ptr = h->theta_part; *chunk_bytes = (int64_t)h->n * (THETA_BATCH + 1) * 8; break;
        case GPIRT_XCHG_ITEMS_PACKED: *dev_ptr = h->items_packed; *chunk_bytes = (int64_t)h->m_chunk * (h->C + 5) * 8; break;
        default: return fail(GPIRT_ERR_INVALID, "exchange_buffer: unknown buffer %d", which);
    }
    return GPIRT_OK;
}

// One process, several shards (handles of one chain, possibly on different GPUs): chunk `src->rank` of exchange buffer `which` goes from src's
// buffer into dst's, on dst's stream, after everything queued on src's stream so far. A peer copy between two GPUs (direct over NVLink once
// peer access is enabled, gpirt_mcmc does that), a device copy on one. The caller keeps src from rewriting its chunk before dst has pulled it
// (gpirt_mcmc's loop: every round ends in gpirt_theta_items_decide, which synchronises each shard's stream).
int gpirt_exchange_pull(gpirt_sampler* dst, gpirt_sampler* src, int which) {
    CHECK_H(dst);
    CHECK_H(src);
    if (dst == src) return GPIRT_OK;
    if (dst->world != src->world || dst->n != src->n || dst->m_total != src->m_total || dst->C != src->C)
        return fail(GPIRT_ERR_SHAPE, "exchange_pull: the two handles are not shards of the same problem");
    void *pd = nullptr, *ps = nullptr;
    int64_t bd = 0, bs = 0;
    TRY(gpirt_exchange_buffer(dst, which, &pd, &bd));
    TRY(gpirt_exchange_buffer(src, which, &ps, &bs));
    if (bd != bs || !pd || !ps) return fail(GPIRT_ERR_SHAPE, "exchange_pull: buffer %d differs between the shards", which);
    const size_t off = (size_t)src->rank * (size_t)bs;
    CU(cudaSetDevice(src->cfg.device));
    CU(cudaEventRecord(src->ev_xchg, src->stream));
    CU(cudaSetDevice(dst->cfg.device));
    CU(cudaStreamWaitEvent(dst->stream, src->ev_xchg, 0));
    CU(cudaMemcpyPeerAsync((char*)pd + off, dst->cfg.device, (char*)ps + off, src->cfg.device, (size_t)bs, dst->stream));
    return GPIRT_OK;
}

// ---- stateless leaves ----
int gpirt_K(int device, const double* x1, int64_t n1, const double* x2, int64_t n2, const double* sds3, double* out) {
    if (!x1 || !x2 || !sds3 || !out || n1 < 1 || n2 < 1) return fail(GPIRT_ERR_INVALID, "gpirt_K: bad arguments");
    TRY(select_device(device));
    DevTemp t1, t2, tK;
    CU(t1.alloc(n1 * 8)); CU(t2.alloc(n2 * 8)); CU(tK.alloc(n1 * n2 * 8));
    double *d1 = t1.as<double>(), *d2 = t2.as<double>(), *dK = tK.as<double>();
    CU(cudaMemcpy(d1, x1, n1 * 8, cudaMemcpyHostToDevice)); CU(cudaMemcpy(d2, x2, n2 * 8, cudaMemcpyHostToDevice));
    CU(build_K(0, d1, (int)n1, d2, (int)n2, sds3, dK, (int)n1, 0.0));
    CU(cudaMemcpy(out, dK, n1 * n2 * 8, cudaMemcpyDeviceToHost));
    return GPIRT_OK;
}
int gpirt_K_time(int device, const double* x1, int64_t n1, const double* x2, int64_t n2, double os, double ls, const double* tsds2,
                 int kernel, double* out) {
    if (!x1 || !x2 || !tsds2 || !out || n1 < 1 || n2 < 1) return fail(GPIRT_ERR_INVALID, "gpirt_K_time: bad arguments");
    TRY(select_device(device));
    DevTemp t1, t2, tK;
    CU(t1.alloc(n1 * 8)); CU(t2.alloc(n2 * 8)); CU(tK.alloc(n1 * n2 * 8));
    double *d1 = t1.as<double>(), *d2 = t2.as<double>(), *dK = tK.as<double>();
    CU(cudaMemcpy(d1, x1, n1 * 8, cudaMemcpyHostToDevice)); CU(cudaMemcpy(d2, x2, n2 * 8, cudaMemcpyHostToDevice));
    CU(cudaMemset(dK, 0, n1 * n2 * 8));  // an unknown KERNEL string leaves the matrix untouched in the reference; here zeros
    CU(build_K_time(0, d1, (int)n1, d2, (int)n2, os, ls, tsds2[0], tsds2[1], kernel, dK, (int)n1));
    CU(cudaMemcpy(out, dK, n1 * n2 * 8, cudaMemcpyDeviceToHost));
    return GPIRT_OK;
}
int gpirt_chol(int device, const double* A, int64_t n, double* L) {
    if (!A || !L || n < 1) return fail(GPIRT_ERR_INVALID, "gpirt_chol: bad arguments");
    TRY(select_device(device));
    const int ld = pad_ld(n);
    DevTemp tA, tI;
    int hinfo = 0;
    CU(tA.alloc((size_t)ld * n * 8)); CU(tI.alloc(4 * (1 + chol_ws_ints((int)n))));
    double* dA = tA.as<double>();
    int* info = tI.as<int>();
    CU(cudaMemset(dA, 0, (size_t)ld * n * 8)); CU(cudaMemset(info, 0, 4));
    CU(cudaMemcpy2D(dA, (size_t)ld * 8, A, n * 8, n * 8, n, cudaMemcpyHostToDevice));
    DevTemp tT, tS;
    CholPlan plan;
    if (n < 4096) {
        const std::vector<uint32_t> tasks = chol_plan_tasks((int)n);
        CU(tT.alloc(std::max<size_t>(tasks.size(), 1) * 4));
        CU(tS.alloc(chol_plan_state_ints((int)n) * 4));
        if (!tasks.empty()) CU(cudaMemcpy(tT.p, tasks.data(), tasks.size() * 4, cudaMemcpyHostToDevice));
        plan.tasks = tT.as<uint32_t>(); plan.n_tasks = (int)tasks.size(); plan.state = tS.as<int>(); plan.max_ctas = getenv("GPIRT_CHOL_CTAS") ? atoi(getenv("GPIRT_CHOL_CTAS")) : 48;
    }
    CU(chol_lower_blocked(0, dA, ld, (int)n, info, info + 1, nullptr, nullptr, plan.state ? &plan : nullptr));
    CU(cudaMemcpy2D(L, n * 8, dA, (size_t)ld * 8, n * 8, n, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(&hinfo, info, 4, cudaMemcpyDeviceToHost));
    if (hinfo < 0) return fail(GPIRT_ERR_CUDA, "chol(): the persistent factorisation kernel gave up waiting for a tile");
    if (hinfo) return fail(GPIRT_ERR_NOT_PD, "chol(): decomposition failed at leading minor %d", hinfo);
    return GPIRT_OK;
}
int gpirt_double_solve(int device, const double* L, int64_t n, const double* X, int64_t nrhs, double* out) {
    if (!L || !X || !out || n < 1 || nrhs < 1) return fail(GPIRT_ERR_INVALID, "gpirt_double_solve: bad arguments");
    TRY(select_device(device));
    const int ld = pad_ld(n);
    DevTemp tL, tX;
    CU(tL.alloc((size_t)ld * n * 8)); CU(tX.alloc((size_t)ld * nrhs * 8));
    double *dL = tL.as<double>(), *dX = tX.as<double>();
    CU(cudaMemset(dL, 0, (size_t)ld * n * 8)); CU(cudaMemset(dX, 0, (size_t)ld * nrhs * 8));
    CU(cudaMemcpy2D(dL, (size_t)ld * 8, L, n * 8, n * 8, n, cudaMemcpyHostToDevice));
    CU(cudaMemcpy2D(dX, (size_t)ld * 8, X, n * 8, n * 8, nrhs, cudaMemcpyHostToDevice));
    CU(trsm_left_lower(0, dL, ld, (int)n, dX, ld, (int)nrhs));
    CU(trsm_left_lower_t(0, dL, ld, (int)n, dX, ld, (int)nrhs));
    CU(cudaMemcpy2D(out, n * 8, dX, (size_t)ld * 8, n * 8, nrhs, cudaMemcpyDeviceToHost));
    return GPIRT_OK;
}

}  // extern "C"

extern "C" {

// Initial values the reference draws with R's RNG inside gpirtMCMC (src/gpirtMCMC.cpp:149-163):
// beta0 ~ N(prior mean, prior sd) per coefficient, f0_j = L z_j. Here from the Philox INIT streams.
// f_j = L z_j for the local items, z in zbuf (rmvnorm(L), src/mvnormal.h:7-14)
static int f_from_zbuf(gpirt_sampler* h) {
    GemmArgs g{};
    g.M = h->n; g.N = h->m_local; g.K = h->n;
    g.A = h->L; g.lda = h->ldn; g.B = h->zbuf; g.ldb = h->ldz; g.C = h->f; g.ldc = h->ldn;
    g.alpha = 1.0; g.beta = 0.0; g.a_lower = 1;
    CU(gemm_f64(h->stream, false, false, g));
    return GPIRT_OK;
}

int gpirt_init_draw(gpirt_sampler* h, const double* beta_prior_means, int draw_beta, int draw_f) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    cudaStream_t st = h->stream;
    if (draw_beta) {
        if (h->cfg.p_beta != 3) return fail(GPIRT_ERR_SHAPE, "init_draw: beta has 3 rows in gpirtMCMC");
        DevTemp t_means;
        CU(t_means.alloc(3 * (size_t)h->m_total * 8));
        double* d_means = t_means.as<double>();
        if (beta_prior_means) CU(cudaMemcpyAsync(d_means, beta_prior_means, 3 * (size_t)h->m_total * 8, cudaMemcpyHostToDevice, st));
        else CU(cudaMemsetAsync(d_means, 0, 3 * (size_t)h->m_total * 8, st));
        k_init_beta<<<(h->m_total + 127) / 128, 128, 0, st>>>(h->beta, d_means, h->sds, h->m_total, h->key());
        ++g_kernel_launches;
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(st));
    }
    if (draw_f && h->m_local > 0) {
        if (!h->chol_valid) TRY(do_update_cholesky(h));
        TRY(wait_L(h));
        CU(launch_gen_normals(st, h->zbuf, h->ldz, h->n, h->m_local, h->key(), ST_INIT_F, h->item0));
        TRY(f_from_zbuf(h));
    }
    return GPIRT_OK;
}

int gpirt_init_f_from_z(gpirt_sampler* h, const double* z) {
    CHECK_H(h);
    if (!z) return fail(GPIRT_ERR_INVALID, "init_f_from_z: NULL z");
    CU(cudaSetDevice(h->cfg.device));
    if (h->m_local <= 0) return GPIRT_OK;
    if (!h->chol_valid) TRY(do_update_cholesky(h));
    TRY(wait_L(h));
    CU(cudaMemcpy2DAsync(h->zbuf, (size_t)h->ldz * 8, z + (size_t)h->item0 * h->n, (size_t)h->n * 8, (size_t)h->n * 8, h->m_local,
                         cudaMemcpyHostToDevice, h->stream));
    TRY(f_from_zbuf(h));
    CU(cudaStreamSynchronize(h->stream));  // z is the caller's memory
    return GPIRT_OK;
}


// ================================================================================================
// horizon > 1 (dynamic GPIRT), constant_IRF == 0. Given theta the sessions are independent (the reference loops h outside every
// per-item block: src/draw-f.cpp:62, src/draw-fstar.cpp:20, src/draw-beta.cpp:60, src/draw_threshold.cpp:206; gather and K per session:
// src/gpirtMCMC.cpp:291-298, src/cholesky-cache.cpp:20-24), so a panel is H single-session samplers on one main stream (each keeps its
// own side stream for its theta-only chain) plus the joint theta step of src/draw-theta.cpp:171-236.
struct gpirt_panel {
    gpirt_config cfg;
    int H = 0, regime = 0;  // 0: H == 1 (the session's own rule); 1: RDM (ls <= 0.1); 2: GP over time (0.1 < ls < 3H)
    std::vector<gpirt_sampler*> ses;
    std::vector<double> L_time;      // H x H, host
    double* d_L_time = nullptr;
    ThetaSession* d_ses = nullptr;
};

static int panel_fail_free(gpirt_panel* p, int rc) {
    gpirt_panel_destroy(p);
    return rc;
}

int gpirt_panel_create(const gpirt_config* cfg, gpirt_panel** out) {
    if (!cfg || !out) return fail(GPIRT_ERR_INVALID, "gpirt_panel_create: NULL argument");
    *out = nullptr;
    const int H = cfg->horizon;
    if (H < 1 || H > GPIRT_MAX_HORIZON) return fail(GPIRT_ERR_SHAPE, "horizon %d outside 1..%d", H, GPIRT_MAX_HORIZON);
    if (H > 1 && cfg->world != 1) return fail(GPIRT_ERR_UNSUPPORTED, "horizon > 1 on an item shard: not built");
    if (H > 1 && cfg->theta_ls >= 3.0 * H)
        return fail(GPIRT_ERR_UNSUPPORTED, "theta_ls = %g >= 3 * horizon (CST regime) with horizon %d: the reference indexes thresholds(j, c) with j up to "
                    "m*horizon-1 on an m-row matrix there (src/draw-theta.cpp:139,162 -> :17-19) and aborts; use theta_ls <= 0.1 (RDM) or "
                    "0.1 < theta_ls < 3*horizon (GP over time)", cfg->theta_ls, H);
    gpirt_panel* p = new gpirt_panel();
    p->cfg = *cfg; p->H = H;
    p->regime = H == 1 ? 0 : (cfg->theta_ls <= 0.1 ? 1 : 2);
    for (int h = 0; h < H; ++h) {
        gpirt_config c = *cfg;
        c.horizon = 1;
        if (p->regime == 1) c.theta_ls = 0.0;  // the session's own theta step uses L_i = sqrt(1 + sd0_i^2) (src/draw-theta.cpp:194)
        gpirt_sampler* s = nullptr;
        int rc = gpirt_create(&c, &s);
        if (s) p->ses.push_back(s);
        if (rc != GPIRT_OK) return panel_fail_free(p, rc);
        s->session = h;
        if (h > 0) {
            rc = gpirt_set_stream(s, (void*)p->ses[0]->stream);
            if (rc != GPIRT_OK) return panel_fail_free(p, rc);
            // QUIRK kept: the RDM branch scores every session's theta against SESSION 0's thresholds (the whole cube goes to a 1-session
            // ESS that reads slice 0: src/draw-theta.cpp:196-198 -> :44-47)
            if (p->regime == 1) s->thr_theta = p->ses[0]->thr;
        }
    }
    if (p->regime == 2) {
        // L_time = chol(K_time(ts, ts, os, ls, (0, 0), KERNEL) + 1e-6 I), ts = 0..H-1; the same factor for every respondent
        // (src/cholesky-cache.cpp:27-35 passes zeros for the linear terms). Host arithmetic, H is tiny.
        std::vector<double>& L = p->L_time;
        L.assign((size_t)H * H, 0.0);
        const double os = cfg->theta_os, ls = cfg->theta_ls;
        for (int j = 0; j < H; ++j)
            for (int i = 0; i < H; ++i) {
                const double diff = std::fabs((double)i - (double)j);
                double r;
                if (cfg->kernel == 0) {  // src/covariance-function.cpp:31-35
                    r = os * os * (1 + std::sqrt(5.0) * diff / ls + 5 * diff * diff / ls / ls / 3);
                    r = r * std::exp(-std::sqrt(5.0) * diff / ls);
                } else {                 // :36-38
                    r = os * os * std::exp(-diff * diff / ls / ls);
                }
                r += (double)i * 0.0 * (double)j;
                r += 0.0;
                L[i + (size_t)j * H] = r + (i == j ? 1e-6 : 0.0);
            }
        for (int j = 0; j < H; ++j) {  // left-looking column Cholesky, upper triangle zeroed (arma::chol(V, "lower"))
            double* cj = &L[(size_t)j * H];
            for (int k = 0; k < j; ++k) {
                const double* ck = &L[(size_t)k * H];
                const double ljk = ck[j];
                if (ljk == 0.0) continue;
                for (int i = j; i < H; ++i) cj[i] -= ck[i] * ljk;
            }
            if (!(cj[j] > 0.0)) { fail(GPIRT_ERR_NOT_PD, "chol(): decomposition failed: K_time + 1e-6 I not positive definite at leading minor %d", j + 1); return panel_fail_free(p, GPIRT_ERR_NOT_PD); }
            const double d = std::sqrt(cj[j]);
            cj[j] = d;
            for (int i = j + 1; i < H; ++i) cj[i] /= d;
            for (int i = 0; i < j; ++i) cj[i] = 0.0;
        }
        if (dev_malloc((void**)&p->d_L_time, (size_t)H * H * 8) != cudaSuccess || dev_malloc((void**)&p->d_ses, (size_t)H * sizeof(ThetaSession)) != cudaSuccess ||
            cudaMemcpy(p->d_L_time, L.data(), (size_t)H * H * 8, cudaMemcpyHostToDevice) != cudaSuccess) {
            fail(GPIRT_ERR_CUDA, "panel: device allocation failed");
            return panel_fail_free(p, GPIRT_ERR_CUDA);
        }
    }
    *out = p;
    return GPIRT_OK;
}

int gpirt_panel_destroy(gpirt_panel* p) {
    if (!p) return GPIRT_OK;
    for (gpirt_sampler* s : p->ses) {  // sessions 1.. borrowed session 0's stream: give them none to destroy
        if (s != p->ses[0]) { cudaSetDevice(s->cfg.device); quiesce(s); cudaStreamSynchronize(s->stream); s->own_stream = false; }
    }
    for (size_t h = p->ses.size(); h-- > 0;) gpirt_destroy(p->ses[h]);
    if (p->d_L_time) dev_free(p->d_L_time);
    if (p->d_ses) dev_free(p->d_ses);
    delete p;
    return GPIRT_OK;
}

int gpirt_panel_session(gpirt_panel* p, int h, gpirt_sampler** out) {
    if (!p || !out || h < 0 || h >= p->H) return fail(GPIRT_ERR_INVALID, "panel_session: bad arguments");
    *out = p->ses[h];
    return GPIRT_OK;
}

int gpirt_panel_get_L_time(gpirt_panel* p, double* out) {
    if (!p || !out) return fail(GPIRT_ERR_INVALID, "panel_get_L_time: NULL argument");
    if (p->regime != 2) return fail(GPIRT_ERR_INVALID, "L_time exists in the GP-over-time regime only (0.1 < theta_ls < 3*horizon, horizon > 1)");
    std::memcpy(out, p->L_time.data(), p->L_time.size() * 8);
    return GPIRT_OK;
}

// RDM: `s` is an array of H stream sets (or NULL), session h draws with s[h]; GP: one set with H normals per respondent.
int gpirt_panel_draw_theta(gpirt_panel* p, const gpirt_streams* s) {
    if (!p) return fail(GPIRT_ERR_INVALID, "panel_draw_theta: NULL panel");
    CU(cudaSetDevice(p->cfg.device));
    const int H = p->H;
    if (p->regime != 2) {
        for (int h = 0; h < H; ++h) {
            gpirt_sampler* S = p->ses[h];
            TRY(need_data(S));
            TRY(do_draw_theta(S, s ? &s[h] : nullptr));
        }
        for (int h = 0; h < H; ++h) TRY(commit_theta(p->ses[h]));
        return GPIRT_OK;
    }
    gpirt_sampler* S0 = p->ses[0];
    std::vector<ThetaSession> hs(H);
    for (int h = 0; h < H; ++h) {
        gpirt_sampler* S = p->ses[h];
        TRY(need_data(S));
        const double* table = nullptr;
        TRY(do_draw_theta(S, nullptr, &table));  // builds the session's term table when it pays
        hs[h] = ThetaSession{S->y_rm, S->theta, S->theta_x, S->fstar, S->beta, S->thr, table};
    }
    CU(cudaMemcpyAsync(p->d_ses, hs.data(), (size_t)H * sizeof(ThetaSession), cudaMemcpyHostToDevice, S0->stream));
    ThetaPanelView v{};
    v.n = S0->n; v.m = S0->m_total; v.ld_chunk = S0->ld_chunk; v.C = S0->C; v.H = H; v.ldm = S0->ldm;
    v.ses = p->d_ses; v.L_time = p->d_L_time; v.key = S0->key();
    v.evals = S0->evals[GPIRT_BLOCK_THETA]; v.margin = S0->margin[GPIRT_BLOCK_THETA]; v.flags = S0->flags;
    TRY(stage_streams(S0, s, S0->n, (size_t)H, &v.z, &v.z_stride, &v.us));
    SEC(S0, "ess_theta", CU(launch_ess_theta_panel(S0->stream, v)));
    CU(cudaStreamSynchronize(S0->stream));  // hs (host) was the source of an async copy
    for (int h = 0; h < H; ++h) TRY(commit_theta(p->ses[h]));
    return GPIRT_OK;
}

int gpirt_panel_sweep(gpirt_panel* p, int n_sweeps) {
    if (!p) return fail(GPIRT_ERR_INVALID, "panel_sweep: NULL panel");
    CU(cudaSetDevice(p->cfg.device));
    for (gpirt_sampler* S : p->ses) {
        TRY(need_data(S));
        if (S->cfg.p_beta != 3) return fail(GPIRT_ERR_SHAPE, "a sweep needs 3 mean coefficients per item");
    }
    for (int k = 0; k < n_sweeps; ++k) {
        for (gpirt_sampler* S : p->ses) {  // draw_f, draw_fstar (src/gpirtMCMC.cpp:270-277)
            TRY(do_draw_f(S, nullptr));
            TRY(do_draw_fstar(S, nullptr));
        }
        TRY(gpirt_panel_draw_theta(p, nullptr));  // :280-284, commits theta
        for (gpirt_sampler* S : p->ses) {  // :286-331; commit_theta already ran
            TRY(launch_theta_chain(S));
            TRY(do_gather_f(S));
            TRY(do_small_ess(S, IB_BETA, nullptr));
            if (!S->chol_valid) TRY(do_update_cholesky(S));
            TRY(do_small_ess(S, IB_THR, nullptr));
            TRY(do_item_ll(S));
            S->iter += 1;
        }
    }
    return GPIRT_OK;
}

int gpirt_panel_total_ll(gpirt_panel* p, double* total) {
    if (!p || !total) return fail(GPIRT_ERR_INVALID, "panel_total_ll: NULL argument");
    double acc = 0.0;
    for (gpirt_sampler* S : p->ses) {  // one running sum over (h, j) in the reference (src/gpirtMCMC.cpp:324-331); per session, then in session order here
        double t = 0.0;
        TRY(gpirt_total_ll(S, &t));
        acc += t;
    }
    *total = acc;
    return GPIRT_OK;
}

int gpirt_panel_sync(gpirt_panel* p) {
    if (!p) return fail(GPIRT_ERR_INVALID, "panel_sync: NULL panel");
    for (gpirt_sampler* S : p->ses) TRY(gpirt_sync(S));
    return GPIRT_OK;
}

}  // extern "C"

extern "C" {

// Per-section device timing for bench.py's roofline block: CUDA events around each kernel group.
int gpirt_profile_enable(gpirt_sampler* h, int on) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    TRY(quiesce(h));  // profiled sweeps (level 1) run every kernel on the main stream, in the reference's order
    h->profiling = on != 0;
    h->profile_level = on;
    return GPIRT_OK;
}
// joins the side streams without recording the wait as a section (it would show up as a stall of the last sweep that the schedule itself does not have)
static int quiesce_unrecorded(gpirt_sampler* h) {
    const bool was = h->profiling;
    h->profiling = false;
    const int rc = quiesce(h);
    h->profiling = was;
    return rc;
}
// Collects finished sections; writes up to `cap` entries "name\0" (32 bytes each) with total ms and launch counts.
int gpirt_profile_read(gpirt_sampler* h, char* names, double* ms, int64_t* launches, int cap, int* count) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    TRY(quiesce_unrecorded(h));  // sections of the overlapped schedule (profiling level 2) live on the side stream
    CU(cudaStreamSynchronize(h->stream));
    for (auto& p : h->pending) {
        float t = 0.f;
        cudaEventElapsedTime(&t, p.a, p.b);
        auto& s = h->sections[p.name];
        s.first += t; s.second += 1;
        cudaEventDestroy(p.a); cudaEventDestroy(p.b);
    }
    h->pending.clear();
    int k = 0;
    for (auto& kv : h->sections) {
        if (k >= cap) break;
        std::snprintf(names + 32 * k, 32, "%s", kv.first.c_str());
        ms[k] = kv.second.first; launches[k] = kv.second.second;
        ++k;
    }
    *count = k;
    return GPIRT_OK;
}
// Every section recorded since the last read / reset, in the order it was issued, with its begin and end on the device clock relative to the begin of
// the first one (events of different streams of one device share a clock): the overlapped schedule as it ran. Collects like gpirt_profile_read.
int gpirt_profile_timeline(gpirt_sampler* h, char* names, double* begin_ms, double* end_ms, int cap, int* count) {
    CHECK_H(h);
    if (!names || !begin_ms || !end_ms || !count) return fail(GPIRT_ERR_INVALID, "profile_timeline: NULL output");
    CU(cudaSetDevice(h->cfg.device));
    TRY(quiesce_unrecorded(h));
    CU(cudaStreamSynchronize(h->stream));
    int k = 0;
    for (auto& p : h->pending) {
        float t = 0.f, b = 0.f, e = 0.f;
        cudaEventElapsedTime(&t, p.a, p.b);
        if (k < cap) {
            cudaEventElapsedTime(&b, h->pending[0].a, p.a);
            cudaEventElapsedTime(&e, h->pending[0].a, p.b);
            std::snprintf(names + 32 * k, 32, "%s", p.name);
            begin_ms[k] = b; end_ms[k] = e;
            ++k;
        }
        auto& sct = h->sections[p.name];
        sct.first += t; sct.second += 1;
    }
    for (auto& p : h->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    h->pending.clear();
    *count = k;
    return GPIRT_OK;
}
int gpirt_profile_reset(gpirt_sampler* h) {  // forgets the totals AND the sections recorded so far whose events have not been collected yet
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    TRY(quiesce_unrecorded(h));
    CU(cudaStreamSynchronize(h->stream));
    for (auto& p : h->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    h->pending.clear();
    h->sections.clear();
    return GPIRT_OK;
}

// Test hooks: the raw Philox streams the kernels consume, so tests can pin them bit-for-bit against the oracle's restatement.
int gpirt_debug_uniforms(gpirt_sampler* h, uint32_t stream, uint32_t unit, int count, double* out) {
    CHECK_H(h);
    RngKey key = h->key();
    for (int k = 0; k < count; k += 2) {
        double a, b;
        philox_uniform2(key, stream, unit, (uint32_t)(k >> 1), a, b);
        out[k] = a;
        if (k + 1 < count) out[k + 1] = b;
    }
    return GPIRT_OK;
}
int gpirt_debug_normals(gpirt_sampler* h, uint32_t stream, uint32_t unit0, int rows, int cols, double* out) {
    CHECK_H(h);
    CU(cudaSetDevice(h->cfg.device));
    DevTemp t_d;
    CU(t_d.alloc((size_t)rows * cols * 8));
    double* d = t_d.as<double>();
    CU(launch_gen_normals(h->stream, d, rows, rows, cols, h->key(), stream, unit0));
    CU(cudaMemcpyAsync(out, d, (size_t)rows * cols * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return GPIRT_OK;
}

}  // extern "C"

extern "C" long long gpirt_launch_count(void) { return gpirt::g_kernel_launches; }

// Test hook: the device Phi / log the likelihood kernels use, evaluated element-wise (tests pin them against mpmath).
__global__ void k_debug_probit(const double* x, int n, double* phi_out, double* log_out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    phi_out[i] = phi(x[i]);
    log_out[i] = x[i] > 0.0 ? log_pos(x[i]) : 0.0;
}
extern "C" int gpirt_debug_probit(int device, const double* x, int n, double* phi_out, double* log_out) {
    if (!x || !phi_out || !log_out || n < 1) return fail(GPIRT_ERR_INVALID, "debug_probit: bad arguments");
    TRY(select_device(device));
    DevTemp t_d;
    CU(t_d.alloc((size_t)3 * n * 8));
    double* d = t_d.as<double>();
    CU(cudaMemcpy(d, x, (size_t)n * 8, cudaMemcpyHostToDevice));
    k_debug_probit<<<(n + 255) / 256, 256>>>(d, n, d + n, d + 2 * (size_t)n);
    CU(cudaGetLastError());
    CU(cudaMemcpy(phi_out, d + n, (size_t)n * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(log_out, d + 2 * (size_t)n, (size_t)n * 8, cudaMemcpyDeviceToHost));
    return GPIRT_OK;
}

// Test hook: the fp32 screening term, its error bound, and the fp64 cell term the ESS kernels use, element-wise.
__global__ void k_debug_cell32(const double* g, const double* lo, const double* hi, int n, double* t32, double* e32, double* t64) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float t, e;
    cell_ll32(g[i], lo[i], hi[i], t, e);
    // the packed two-cell form the ESS kernels run (cell i with its neighbour): the same term bit for bit, the bound up to the rounding of its own sum
    const int k = (i ^ 1) < n ? (i ^ 1) : i;
    float ta, ea, tb, eb;
    if (i & 1) cell_ll32x2(g[k], lo[k], hi[k], g[i], lo[i], hi[i], tb, eb, ta, ea);
    else cell_ll32x2(g[i], lo[i], hi[i], g[k], lo[k], hi[k], ta, ea, tb, eb);
    const bool same = (__float_as_uint(ta) == __float_as_uint(t) || (ta != ta && t != t)) && (fabsf(ea - e) <= 1e-5f * fabsf(e) || (ea != ea && e != e));
    if (!same) e = CUDART_NAN_F;  // the test requires finite bounds
    t32[i] = (double)t;
    e32[i] = (double)fmaxf(e, ea);
    t64[i] = cell_ll(g[i], lo[i], hi[i], isinf(lo[i]) && lo[i] < 0, isinf(hi[i]) && hi[i] > 0);
}
extern "C" int gpirt_debug_cell32(int device, const double* g, const double* tau_lo, const double* tau_hi, int n, double* t32, double* e32,
                                  double* t64) {
    if (!g || !tau_lo || !tau_hi || !t32 || !e32 || !t64 || n < 1) return fail(GPIRT_ERR_INVALID, "debug_cell32: bad arguments");
    TRY(select_device(device));
    DevTemp t_d;
    CU(t_d.alloc((size_t)6 * n * 8));
    double* d = t_d.as<double>();
    CU(cudaMemcpy(d, g, (size_t)n * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d + n, tau_lo, (size_t)n * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d + 2 * (size_t)n, tau_hi, (size_t)n * 8, cudaMemcpyHostToDevice));
    k_debug_cell32<<<(n + 255) / 256, 256>>>(d, d + n, d + 2 * (size_t)n, n, d + 3 * (size_t)n, d + 4 * (size_t)n, d + 5 * (size_t)n);
    CU(cudaGetLastError());
    CU(cudaMemcpy(t32, d + 3 * (size_t)n, (size_t)n * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(e32, d + 4 * (size_t)n, (size_t)n * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(t64, d + 5 * (size_t)n, (size_t)n * 8, cudaMemcpyDeviceToHost));
    return GPIRT_OK;
}

// Measurement hook: fp64 issue-rate peaks of this device (include/gpirt_b200.h). Eight independent accumulator chains per thread, 2048 threads per SM.
__global__ void k_peak_dmma(double* out, int iters, double a, double b) {
    double c[8][2];
    for (int k = 0; k < 8; ++k) { c[k][0] = threadIdx.x; c[k][1] = k; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) dmma884(c[k][0], c[k][1], a, b);
    }
    double s = 0;
    for (int k = 0; k < 8; ++k) s += c[k][0] + c[k][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_peak_dfma(double* out, int iters, double a, double b) {
    double c[8];
    for (int k = 0; k < 8; ++k) c[k] = threadIdx.x + k;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) c[k] = fma(c[k], a, b);
    }
    double s = 0;
    for (int k = 0; k < 8; ++k) s += c[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
extern "C" int gpirt_debug_fp64_peak(int device, double* dmma_tflops, double* dfma_tflops) {
    if (!dmma_tflops || !dfma_tflops) return fail(GPIRT_ERR_INVALID, "fp64_peak: NULL output");
    TRY(select_device(device));
    int sms = 0;
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int threads = 512, blocks = sms * 4, iters = 20000;
    DevTemp t_out;
    CU(t_out.alloc((size_t)blocks * threads * 8));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    float ms = 0.f;
    double best[2] = {0.0, 0.0};
    for (int rep = 0; rep < 3; ++rep) {  // first repetition warms the clocks up; best of three
        for (int which = 0; which < 2; ++which) {
            cudaEventRecord(e0, 0);
            if (which == 0) k_peak_dmma<<<blocks, threads>>>(t_out.as<double>(), iters, 1.0000001, 1e-9);
            else k_peak_dfma<<<blocks, threads>>>(t_out.as<double>(), iters, 1.0000001, 1e-9);
            cudaEventRecord(e1, 0);
            cudaError_t e = cudaEventSynchronize(e1);
            if (e != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return fail(GPIRT_ERR_CUDA, "fp64_peak: %s", cudaGetErrorString(e)); }
            cudaEventElapsedTime(&ms, e0, e1);
            const double flop = which == 0 ? 2.0 * 256 * 8 * iters * (double)blocks * (threads / 32) : 2.0 * 8 * iters * (double)blocks * threads;
            best[which] = std::max(best[which], flop / ms / 1e9);
        }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *dmma_tflops = best[0];
    *dfma_tflops = best[1];
    return GPIRT_OK;
}

// Measurement hook: the persistent Cholesky with clock64 stamps of its critical CTA (16 per panel step) and of its first worker (8 per task, 256 tasks);
// trace_out holds 16 T + 8 * 256 values (T = ceil(n / 64)). tools/chol_trace.py turns them into a per-phase table.
extern "C" int gpirt_debug_chol_trace(int device, const double* A, int64_t n, int ctas, long long* trace_out) {
    if (!A || !trace_out || n < 1 || n >= 4096) return fail(GPIRT_ERR_INVALID, "chol_trace: bad arguments");
    TRY(select_device(device));
    const int ld = pad_ld(n), T = chol_plan_tiles((int)n);
    const size_t tl = (size_t)16 * T + 8 * 256;
    DevTemp tA, tI, tT, tS, tR;
    CU(tA.alloc((size_t)ld * n * 8)); CU(tI.alloc(4)); CU(tR.alloc(tl * 8));
    const std::vector<uint32_t> tasks = chol_plan_tasks((int)n);
    CU(tT.alloc(std::max<size_t>(tasks.size(), 1) * 4));
    CU(tS.alloc(chol_plan_state_ints((int)n) * 4));
    if (!tasks.empty()) CU(cudaMemcpy(tT.p, tasks.data(), tasks.size() * 4, cudaMemcpyHostToDevice));
    CU(cudaMemset(tA.p, 0, (size_t)ld * n * 8)); CU(cudaMemset(tI.p, 0, 4)); CU(cudaMemset(tR.p, 0, tl * 8));
    CU(cudaMemcpy2D(tA.p, (size_t)ld * 8, A, n * 8, n * 8, n, cudaMemcpyHostToDevice));
    CU(chol_lower_persistent(0, tA.as<double>(), ld, (int)n, tI.as<int>(), tS.as<int>(), tT.as<uint32_t>(), (int)tasks.size(), ctas, tR.as<long long>()));
    CU(cudaMemcpy(trace_out, tR.p, tl * 8, cudaMemcpyDeviceToHost));
    return GPIRT_OK;
}

// Test hook: phi_tab(), the table-based Phi of the log-likelihood pass (probit.cuh), evaluated element-wise from the device copy of its table.
__global__ void k_debug_phi_tab(const double* x, int n, double* out) {
    __shared__ PhiNode s_phi[PHI_TAB_N];
    load_phi_table(s_phi);
    __syncthreads();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = phi_tab(x[i], s_phi);
}
extern "C" int gpirt_debug_phi_tab(int device, const double* x, int n, double* out) {
    if (!x || !out || n < 1) return fail(GPIRT_ERR_INVALID, "debug_phi_tab: bad arguments");
    TRY(select_device(device));
    DevTemp t_d;
    CU(t_d.alloc((size_t)2 * n * 8));
    double* d = t_d.as<double>();
    CU(cudaMemcpy(d, x, (size_t)n * 8, cudaMemcpyHostToDevice));
    k_debug_phi_tab<<<(n + 255) / 256, 256>>>(d, n, d + n);
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, d + n, (size_t)n * 8, cudaMemcpyDeviceToHost));
    return GPIRT_OK;
}
