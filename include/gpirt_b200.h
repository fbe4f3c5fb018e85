/* gpirt_b200 — C-ABI of the B200-native gpirt sweep (drop-in boundary).
 *
 * Every entry point replaces one interface of the reference (paths relative to the reference tree):
 *   gpirt_K                      <- arma::mat K(x1, x2, sds)                    src/gpirt.h:269, src/covariance-function.cpp:3-18
 *   gpirt_K_time                 <- arma::mat K_time(...)                       src/gpirt.h:270-272, src/covariance-function.cpp:21-44
 *   gpirt_delta_to_threshold /
 *   gpirt_threshold_to_delta     <- src/gpirt.h:285-286, src/log-likelihood.cpp:66-88
 *   gpirt_update_cholesky_cache  <- update_cholesky_cache()                     src/gpirt.h:262-266, src/cholesky-cache.cpp:11-40
 *   gpirt_double_solve           <- double_solve(L, X)                          src/gpirt.h:289, src/cholesky-cache.cpp:4-9
 *   gpirt_ll_bar                 <- ll_bar_sparse() over every item             src/gpirt.h:280-282, src/log-likelihood.cpp:50-64,
 *                                                                                src/gpirtMCMC.cpp:324-331
 *   gpirt_draw_f                 <- draw_f()                                    src/gpirt.h:216-222, src/draw-f.cpp:47-83
 *   gpirt_draw_fstar             <- draw_fstar()                                src/gpirt.h:225-232, src/draw-fstar.cpp:4-57
 *   gpirt_draw_theta             <- draw_theta()                                src/gpirt.h:235-244, src/draw-theta.cpp:86-237
 *   gpirt_gather_f               <- f(k,:,h) = f_star(idx,:,h)                  src/gpirtMCMC.cpp:291-298
 *   gpirt_draw_beta              <- draw_beta()                                 src/gpirt.h:247-252, src/draw-beta.cpp:51-101
 *   gpirt_draw_threshold         <- draw_threshold()                            src/gpirt.h:255-259, src/draw_threshold.cpp:171-230
 *   gpirt_sweep                  <- one body of the sweep loop                  src/gpirtMCMC.cpp:261-331
 *   gpirt_mcmc                   <- Rcpp::List gpirtMCMC(...)                   src/gpirtMCMC.cpp:25-406 (.Call _gpirt_gpirtMCMC,
 *                                                                                src/RcppExports.cpp:16-39)
 *   gpirt_recover_fstar          <- Rcpp::List recover_fstar(...)               src/recover_fstar.cpp:8-93 (.Call _gpirt_recover_fstar,
 *                                                                                src/RcppExports.cpp:42-58)
 *   gpirt_create/destroy, set/get <- CholeskyCache + WorkspacePool lifetime     src/gpirt.h:49-68,153-195, src/gpirtMCMC.cpp:89-90
 *
 * Conventions: every matrix is column-major float64 exactly as Armadillo lays it out (a cube is
 * `horizon` contiguous slices); y is float64 with NaN = missing and integer codes 1..C otherwise
 * (gpirt_set_data_i8 takes the compact form: 0 = missing). All pointers in this header are HOST
 * pointers unless a name says `dev`. No call throws; each returns a gpirt_status and
 * gpirt_last_error() holds the message (R shim: Rcpp::stop(gpirt_last_error())). There is no CPU
 * fallback: without a CUDA device every compute call returns GPIRT_ERR_NO_DEVICE.
 *
 * Scope of this build: a gpirt_sampler is one session with constant_IRF == 0 (all BASELINE configs); horizon > 1 is a gpirt_panel
 * (RDM and GP-over-time theta regimes); constant_IRF == 1 (any horizon) runs through gpirt_mcmc / gpirt_recover_fstar as a composition of those handles.
 * What is not built returns GPIRT_ERR_UNSUPPORTED.
 */
#ifndef GPIRT_B200_H
#define GPIRT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPIRT_NGRID 1001     /* theta* = regspace(-5, 0.01, 5), src/gpirtMCMC.cpp:205 */
#define GPIRT_MAX_C 21       /* Workspace::delta_prime(20), src/gpirt.h:111 */

typedef enum {
    GPIRT_OK = 0,
    GPIRT_ERR_INVALID = 1,        /* NULL handle / argument */
    GPIRT_ERR_SHAPE = 2,          /* inconsistent sizes (Armadillo would throw a shape error) */
    GPIRT_ERR_CATEGORIES = 3,     /* C < 2 or C > 21, or a y code outside 1..C */
    GPIRT_ERR_NOT_PD = 4,         /* chol failed (arma::chol throws); gpirt_last_error names the leading minor */
    GPIRT_ERR_CUDA = 5,
    GPIRT_ERR_UNSUPPORTED = 6,    /* a mode that is not built (e.g. constant_IRF != 0 on a single handle - gpirt_mcmc composes it -, theta_ls >= 3H at H > 1) */
    GPIRT_ERR_STREAM_OVERRUN = 7, /* a supplied uniform stream ran dry before an ESS loop accepted */
    GPIRT_ERR_NO_DEVICE = 8,
    GPIRT_ERR_ESS_STALLED = 9     /* an ESS loop hit the proposal cap (reference would spin forever) */
} gpirt_status;

typedef struct gpirt_sampler gpirt_sampler; /* opaque: owns all device memory, streams and events */

typedef struct {
    int32_t n;             /* respondents */
    int32_t m_total;       /* items of the whole problem */
    int32_t horizon;       /* 1 for gpirt_create; gpirt_panel_create takes horizon >= 1 */
    int32_t C;             /* response categories, 2..21 */
    int32_t constant_IRF;  /* must be 0 */
    int32_t p_beta;        /* rows of beta: 3 (gpirtMCMC) or 2 (recover_fstar, src/recover_fstar.cpp:58-60) */
    double theta_os;       /* gpirtMCMC(theta_os=) */
    double theta_ls;       /* gpirtMCMC(theta_ls=): regime switch of src/draw-theta.cpp:112,171,206 */
    int32_t kernel;        /* 0 "Matern", 1 "RBF" (src/covariance-function.cpp:31,36) */
    int32_t device;        /* CUDA ordinal */
    uint64_t seed;         /* Philox key for production-mode streams */
    uint32_t chain;        /* stream id of this chain */
    /* item shard owned by this handle (single GPU: rank 0 of 1). Items [rank*m_total/world, ...). */
    int32_t rank, world;
} gpirt_config;

/* Random streams for one block call. NULL members => on-device Philox4x32-10 keyed by
 * (seed, chain, iter, block, unit). Supplied members (parity mode) are host arrays:
 *   z  : z[unit * z_stride + i]   standard normals, in the reference's draw order
 *   u  : u[unit * u_stride + k]   U(0,1): k=0 slice height, k=1 initial angle, k>=2 one per rejection
 *   nu : draw_f only — the prior draw nu = L z itself, n x m column-major (skips the L*Z contraction)
 * `unit` is the GLOBAL item (or respondent) index. */
typedef struct {
    const double* z;  int64_t z_stride;
    const double* u;  int64_t u_stride;
    const double* nu;
} gpirt_streams;

typedef enum { GPIRT_BLOCK_F = 0, GPIRT_BLOCK_FSTAR = 1, GPIRT_BLOCK_THETA = 2, GPIRT_BLOCK_BETA = 3, GPIRT_BLOCK_THR = 4 } gpirt_block;

const char* gpirt_last_error(void);
int gpirt_device_count(void);

/* ---- stateless leaves (host in, host out; run on `device`) ---- */
int gpirt_K(int device, const double* x1, int64_t n1, const double* x2, int64_t n2, const double* sds3, double* out);
int gpirt_K_time(int device, const double* x1, int64_t n1, const double* x2, int64_t n2, double os, double ls,
                 const double* tsds2, int kernel, double* out);
int gpirt_delta_to_threshold(const double* deltas, int C, double* thresholds);   /* host-only bookkeeping, bit-exact */
int gpirt_threshold_to_delta(const double* thresholds, int C, double* deltas);
int gpirt_chol(int device, const double* A, int64_t n, double* L);               /* arma::chol(A,"lower") */
/* NO COUNTERPART IN THE REFERENCE. gpirt_chol's factorisation for callers that cannot allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): it calls no cudaMalloc / cudaFree or any
 * allocating routine. DEVICE in, DEVICE out: A_dev is the n x n matrix on `device`, column-major with leading dimension ld (even, >= n rounded up
 * to even; A_dev 16-byte aligned), lower triangle read; on return of the work queued on `cuda_stream` (a cudaStream_t, NULL = default stream)
 * the lower triangle holds L, the strict upper triangle zeros, and *info_dev (a device int the call clears first) 0, or the 1-based index of the
 * failing leading minor, or -1 if the kernel gave up waiting. The call does not synchronise. Limit: n <= GPIRT_NOALLOC_CHOL_MAX_N; beyond it the
 * call fails with GPIRT_ERR_SHAPE (use gpirt_chol or a sampler handle). The kernel's scratch (tile version counters + task list, 28 928 bytes)
 * is a static device array that every GPU which loads the library holds until the process exits, and ALL calls on a device share it: it is
 * cleared / rewritten at the start of each call, so calls that overlap on one GPU (two streams, two host threads) race - keep them on one stream. */
#define GPIRT_NOALLOC_CHOL_MAX_N 2048
int gpirt_chol_device_noalloc(int device, double* A_dev, int64_t ld, int64_t n, int32_t* info_dev, void* cuda_stream);
/* arma::interp1(X, Y.col(j), XI, YI.col(j), "linear") for every column j of Y (src/draw-fstar.cpp:77-82, the inducing points of the constant-IRF
 * f* draw): X sorted, duplicated abscissae dropped keeping the lowest row index, NaN outside [min X, max X]. Y: nx x m, YI: nq x m, column-major. */
int gpirt_interp1(int device, const double* X, int64_t nx, const double* Y, int64_t m, const double* XI, int64_t nq, double* YI);
int gpirt_double_solve(int device, const double* L, int64_t n, const double* X, int64_t nrhs, double* out);

/* ---- lifetime, data, state ---- */
int gpirt_create(const gpirt_config* cfg, gpirt_sampler** out);
int gpirt_destroy(gpirt_sampler* h);
/* Large device blocks of destroyed handles are kept in a per-process cache and handed to the next handle of the same shape (a gpirtMCMC()
 * call creates and destroys one per chain); this returns them to the driver. */
int gpirt_release_cached_memory(void);
int gpirt_set_stream(gpirt_sampler* h, void* cuda_stream);  /* run on the caller's stream (e.g. torch's current stream) */
int gpirt_set_data_f64(gpirt_sampler* h, const double* y);  /* n x m_total, NaN missing */
int gpirt_set_data_i8(gpirt_sampler* h, const int8_t* y);   /* n x m_total, 0 missing */
/* chains share their responses (R/gpirtMCMC.R:111-158 passes the same `data` to every chain): take src's device layouts of y, no upload */
int gpirt_copy_data(gpirt_sampler* h, gpirt_sampler* src);
/* constant_IRF == 1 building block: per-row means for draw_f / draw_threshold on a sampler whose rows are the stacked (session, respondent)
 * pairs (src/draw-f.cpp:93-104 stacks mu.slice(h)); mu is n x m_total (host), NULL switches back to X beta */
int gpirt_set_mean_rows(gpirt_sampler* h, const double* mu);
int gpirt_set_priors(gpirt_sampler* h, const double* beta_prior_sds /*3 x m_total*/, const double* theta_prior_sds /*2 x n*/);
/* any pointer may be NULL (left unchanged). beta: p_beta x m_total; f: n x m_total; thresholds: m_total x (C+1);
 * fstar: NGRID x m_total. Setting theta clamps it to [-5,5] (src/gpirtMCMC.cpp:133). */
int gpirt_set_state(gpirt_sampler* h, const double* theta, const double* beta, const double* f, const double* thresholds,
                    const double* fstar);
int gpirt_get_state(gpirt_sampler* h, double* theta, double* beta, double* f, double* thresholds, double* fstar);
int gpirt_get_L(gpirt_sampler* h, double* L /* n x n, upper triangle zero */);
/* Device-side store of the draws gpirtMCMC() keeps (theta_draws / beta_draws / threshold_draws / ll_draws / f_draws / fstar_draws,
 * src/gpirtMCMC.cpp:230-255,334-356). gpirt_store_push appends the current state as sample number `count` WITHOUT a host
 * synchronisation (copies and the ordered log-likelihood sum are enqueued on the sampler's stream, in the reference's layouts);
 * gpirt_store_flush synchronises once, reports any error raised by the sweeps since the last synchronisation, and writes the
 * `count` buffered samples as samples first .. first+count-1 of the reference's output arrays (theta: cube n_samples x n x 1;
 * per sample: beta 3 x m, thresholds m x (C+1), f n x m, fstar NGRID x m; ll: n_samples), then empties the store.
 * NULL outputs are skipped. On an item shard (world > 1) theta, beta, thresholds and ll can be stored — they are complete on every
 * rank after the sweep's exchanges — f and f* cannot (each rank holds its own items). */
int gpirt_store_open(gpirt_sampler* h, int capacity, int store_f, int store_fstar);
int gpirt_store_push(gpirt_sampler* h);
int gpirt_store_count(gpirt_sampler* h, int* count, int* capacity);
int gpirt_store_flush(gpirt_sampler* h, int n_samples, int first, double* theta_draws, double* beta_draws, double* threshold_draws,
                      double* ll_draws, double* f_draws, double* fstar_draws);
/* the same for session `slice` of a horizon-`horizon` run: outputs are the reference's cubes with `horizon` slices per sample; ll_draws is
 * set by slice 0 and added to by later slices (flush the sessions in order) */
int gpirt_store_flush_slice(gpirt_sampler* h, int n_samples, int first, int horizon, int slice, double* theta_draws, double* beta_draws,
                            double* threshold_draws, double* ll_draws, double* f_draws, double* fstar_draws);
int gpirt_store_close(gpirt_sampler* h);
/* bytes of device memory one stored sample takes (to size `capacity`) */
int64_t gpirt_store_sample_bytes(gpirt_sampler* h, int store_f, int store_fstar);
/* f_j = L z_j for every local item from host normals z (n x m_total): the reference's initial f (src/gpirtMCMC.cpp:158-162) */
int gpirt_init_f_from_z(gpirt_sampler* h, const double* z);
int gpirt_set_iter(gpirt_sampler* h, uint32_t iter);         /* sweep counter used to key the Philox streams */
/* How draw_fstar applies L^-1 / L^-T to k* (solve(trimatl(L), kstar) and double_solve, src/draw-fstar.cpp:30,51, src/cholesky-cache.cpp:4-9):
 * 0 = blocked forward/back substitution (backward stable, n/64 dependent steps); 1 = explicit L^-1 by parallel block inversion, then
 * GEMMs (default; forward error kappa(L)*eps like substitution's bound, no dependent chain). */
int gpirt_set_solver(gpirt_sampler* h, int solver);
/* 1 (default): inside gpirt_sweep*, the theta-only chain (K, Cholesky cache, shared draw_fstar terms: src/cholesky-cache.cpp:20-24,
 * src/draw-fstar.cpp:26-36) of the next sweep runs on a second CUDA stream under the item samplers; 0: every kernel in the
 * reference's order on one stream. Results are identical either way. */
int gpirt_set_overlap(gpirt_sampler* h, int on);
/* The Cholesky cache (src/cholesky-cache.cpp:20-24) and chol(K_post) (src/draw-fstar.cpp:36) run as one persistent kernel each (csrc/chol_persistent.cu)
 * holding at most `ctas` CTAs (default 48, GPIRT_CHOL_CTAS); 0 selects the per-panel kernels of csrc/linalg.cu. Same factor up to rounding order;
 * gpirt_mcmc_chains divides the budget among the chains it runs side by side. */
int gpirt_set_chol_ctas(gpirt_sampler* h, int ctas);
/* 1 (default): the ESS kernels (draw_f / draw_beta / draw_threshold: src/draw-f.cpp:23-41 and its three siblings) first evaluate a
 * proposal's log-likelihood in fp32 together with a rigorous bound on its distance to the fp64 value and only run the fp64 pass
 * when the accept test ll' > log_y is closer than that bound. Decisions, states and stream consumption are identical to 0 (every
 * evaluation in fp64). gpirt_get_exact_evals: per unit, how many evaluations of the last call of `block` ran in fp64. */
int gpirt_set_screen(gpirt_sampler* h, int on);
/* Threads per respondent of the theta kernel: 0 = by the number of local respondents (256, or 1024 on a rank with few respondents), or forced;
 * process-wide. Both sum in the same order, so the choice never changes a chain (tests). */
int gpirt_debug_theta_threads(int threads);
int gpirt_get_exact_evals(gpirt_sampler* h, int block, int32_t* out);
int gpirt_get_diag(gpirt_sampler* h, int block, int32_t* evals, double* margin); /* per unit: 1+proposals, min |ll'-log_y| */
int gpirt_sync(gpirt_sampler* h);                            /* wait for queued work; surfaces deferred errors */
/* initial values the reference draws inside gpirtMCMC (src/gpirtMCMC.cpp:149-163): beta0 ~ N(prior mean, prior sd)
 * (beta_prior_means 3 x m_total or NULL = 0) and f0_j = L z_j, from the Philox INIT streams */
int gpirt_init_draw(gpirt_sampler* h, const double* beta_prior_means, int draw_beta, int draw_f);

/* ---- subsystems (asynchronous on the handle's stream unless they return host data) ---- */
int gpirt_update_cholesky_cache(gpirt_sampler* h);
int gpirt_draw_f(gpirt_sampler* h, const gpirt_streams* s);
int gpirt_draw_fstar(gpirt_sampler* h, const gpirt_streams* s);
int gpirt_draw_theta(gpirt_sampler* h, const gpirt_streams* s);
int gpirt_gather_f(gpirt_sampler* h);
int gpirt_draw_beta(gpirt_sampler* h, const gpirt_streams* s);
int gpirt_draw_threshold(gpirt_sampler* h, const gpirt_streams* s);
int gpirt_ll_bar(gpirt_sampler* h, double* per_item /* m_total or NULL */, double* total /* or NULL */);

/* ---- sweep ---- */
int gpirt_sweep(gpirt_sampler* h, int n_sweeps);             /* world == 1: whole sweeps, Philox streams */
/* item-sharded sweeps (world > 1): the caller performs the exchanges between the phases on the
 * same stream; gpirt_exchange_buffer exposes the device buffers they operate on. */
int gpirt_sweep_phase_items(gpirt_sampler* h);               /* draw_f, draw_fstar on the local items */
int gpirt_sweep_phase_theta(gpirt_sampler* h);               /* theta ESS for the local respondents */
int gpirt_sweep_phase_close(gpirt_sampler* h);               /* gather, beta, mu, cholesky, thresholds, item ll */
typedef enum { GPIRT_XCHG_FSTAR = 0, GPIRT_XCHG_BETA = 1, GPIRT_XCHG_THR = 2, GPIRT_XCHG_THETA = 3, GPIRT_XCHG_ITEM_LL = 4,
               GPIRT_XCHG_THETA_PART = 5, GPIRT_XCHG_ITEMS_PACKED = 6 } gpirt_xchg;
/* Item-sharded theta step (the default of the sharded drivers): draw_theta's per-respondent loop (src/draw-theta.cpp:121-170) distributed over the
 * ranks BY ITEMS instead of by respondents. Every rank keeps all n respondents and scores them against its own items (its own term table of
 * NGRID x m_local x C cells); a round evaluates the next 16 proposals of every unfinished respondent at once - the shrink sequence of an ESS unit
 * under rejections depends on its uniforms only (src/draw-theta.cpp:72-80) - and the ranks exchange n x 17 partial log-likelihoods per round
 * (GPIRT_XCHG_THETA_PART, in-place all-gather) instead of the f* table. Sequence per sweep:
 *     gpirt_theta_items_begin(h, streams or NULL);
 *     do { gpirt_theta_items_partials(h); <all-gather GPIRT_XCHG_THETA_PART>; gpirt_theta_items_decide(h, &n_active); } while (n_active > 0);
 *     gpirt_sweep_phase_close(h);   (the new theta is complete on every rank: no theta exchange)
 * gpirt_theta_items_decide synchronises the stream (the loop ends on its count) and is the same computation on every rank: partials added in rank
 * order, then the reference's strict `ll' > log_y` proposal by proposal. The decisions are the single-GPU kernel's unless a margin falls below fp64
 * summation noise; gpirt_get_diag reports the minimum margin per respondent.
 * gpirt_items_pack(h, 0) / <all-gather GPIRT_XCHG_ITEMS_PACKED> / gpirt_items_pack(h, 1): beta, thresholds and the per-item log-likelihood terms
 * of all ranks in ONE exchange instead of three. */
int gpirt_theta_items_begin(gpirt_sampler* h, const gpirt_streams* s);
int gpirt_theta_items_partials(gpirt_sampler* h);
int gpirt_theta_items_decide(gpirt_sampler* h, int32_t* n_active);
int gpirt_items_pack(gpirt_sampler* h, int unpack);
/* One process driving several shards (handles with the same world, ranks 0..world-1, on the same or on different GPUs): copies chunk src->rank of
 * exchange buffer `which` from src's buffer into dst's on dst's stream, ordered after the work queued on src's stream (an event) - the all-gathers
 * above without a process group: a peer copy over NVLink between two GPUs. This is how gpirt_mcmc(n_devices > 1) exchanges; no torch, no NCCL. */
int gpirt_exchange_pull(gpirt_sampler* dst, gpirt_sampler* src, int which);
/* Distributed theta-only chain (config 4 sizes: n >= 4096, item shards). K, its Cholesky factor, L^-1 and K^-1 k* depend on theta alone and are the same on
 * every rank (src/cholesky-cache.cpp:20-24, src/draw-fstar.cpp:26-36); replicated they are 28 of 62 ms per sweep at n = 10 000 on 8 GPUs. With every peer's
 * arena attached, the trailing updates of the Cholesky are split over the ranks by 256-column blocks (the owner of the next block pushes it to its peers
 * over NVLink, every rank factors the panel), the large merge levels of L^-1 by columns of their blocks, and the columns of L^-1 k* and K^-1 k* are
 * computed by their owner and pushed; every tile is computed by the same arithmetic as on one GPU, so the chain does not change (45 instead of 59 ms).
 * Every shard needs a GPU of its own (the exchange points are spinning wait kernels) and the same profiling level on all ranks. gpirt_chain_export: this handle's IPC handle and buffer offsets (offsets6[5]: the arena
 * address, for a peer in the SAME process); gpirt_chain_attach(h, peer_rank, handle64 or NULL for a same-process peer, offsets6). GPIRT_DIST_CHAIN=0
 * keeps the chain replicated. gpirt_chain_detach (then a barrier) precedes gpirt_destroy on every rank. */
int gpirt_chain_export(gpirt_sampler* h, void* handle64, int64_t* offsets6);
int gpirt_chain_attach(gpirt_sampler* h, int peer_rank, const void* handle64, const int64_t* offsets6);
int gpirt_chain_detach(gpirt_sampler* h);
/* test hook, no device needed: what rank `rank` of `world` computes in the distributed chain at n respondents (layout of `out`: csrc/capi.cu); returns the count (>= 4) or GPIRT_ERR_INVALID */
int gpirt_debug_dist_plan(int n, int world, int rank, int32_t* out, int cap);
/* device pointer of the all-gather buffer `which`; it holds `world` equal chunks of chunk_bytes, chunk r owned by rank r */
int gpirt_exchange_buffer(gpirt_sampler* h, int which, void** dev_ptr, int64_t* chunk_bytes);
int gpirt_total_ll(gpirt_sampler* h, double* total);         /* after the ITEM_LL exchange (or world == 1) */
/* The f* exchange (the all-gather that feeds the theta step) without a collective call: every rank maps its peers' device
 * arenas (one process per GPU: cudaIpc handles, exchanged by the caller, e.g. over torch.distributed) and gpirt_sweep_phase_items' last
 * kernel stores each finished f* value into its own table AND into every peer's table over NVLink, then publishes an epoch flag;
 * gpirt_peer_wait enqueues the wait for all peers' flags on the handle's stream in place of the FSTAR all-gather. The theta all-gather
 * that follows the theta step orders a rank's next push after its peers' reads. Without attached peers the phases behave as before and
 * the caller all-gathers GPIRT_XCHG_FSTAR itself. gpirt_ipc_detach (then a barrier) must precede gpirt_destroy on every rank. */
int gpirt_ipc_export(gpirt_sampler* h, void* handle64 /* 64 bytes out: cudaIpcMemHandle_t */, int64_t* fstar_offset, int64_t* flags_offset);
int gpirt_ipc_attach(gpirt_sampler* h, int peer_rank, const void* handle64, int64_t fstar_offset, int64_t flags_offset); /* the PEER's export */
int gpirt_ipc_detach(gpirt_sampler* h);
int gpirt_peer_wait(gpirt_sampler* h);

/* ---- measurement and test hooks (no reference counterpart) ---- */
int gpirt_profile_enable(gpirt_sampler* h, int on);          /* CUDA events around every kernel group of a sweep */
int gpirt_profile_read(gpirt_sampler* h, char* names /* cap x 32 */, double* ms, int64_t* launches, int cap, int* count);
int gpirt_profile_reset(gpirt_sampler* h);
/* level 2 (the overlapped schedule): every recorded section in issue order with its begin / end in ms after the first one's begin (tools/timeline_c3.py) */
int gpirt_profile_timeline(gpirt_sampler* h, char* names /* cap x 32 */, double* begin_ms, double* end_ms, int cap, int* count);
long long gpirt_launch_count(void);                          /* kernels launched by this library so far */
int gpirt_debug_uniforms(gpirt_sampler* h, uint32_t stream, uint32_t unit, int count, double* out);
int gpirt_debug_probit(int device, const double* x, int n, double* phi_out, double* log_out); /* device Phi(x), log(x) */
int gpirt_debug_normals(gpirt_sampler* h, uint32_t stream, uint32_t unit0, int rows, int cols, double* out /* rows x cols */);
/* fp32 screening term, its error bound and the fp64 cell term log(Phi(tau_hi-g) - Phi(tau_lo-g) + 1e-6), element-wise */
int gpirt_debug_phi_tab(int device, const double* x, int n, double* out); /* device Phi(x) of the log-likelihood pass (table + Taylor step, csrc/probit.cuh) */
int gpirt_debug_cell32(int device, const double* g, const double* tau_lo, const double* tau_hi, int n, double* t32, double* e32, double* t64);
/* Measurement hook (bench.py's roofline denominator; MEASURED_PEAKS.json carries no fp64 figure): issue-rate microbenchmarks of the two fp64
 * pipes the sweep's contractions run on, ~20 ms each on an otherwise idle device: mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4) and scalar DFMA,
 * in TFLOP/s. No reference counterpart. */
int gpirt_debug_fp64_peak(int device, double* dmma_tflops, double* dfma_tflops);
/* Measurement hook: the persistent Cholesky (csrc/chol_persistent.cu) on the n x n matrix A (n < 4096) with `ctas` CTAs and clock64 stamps of its
 * critical CTA (16 per panel step) and first worker (8 per task, 256 tasks): trace_out holds 16 ceil(n/64) + 2048 values. */
int gpirt_debug_chol_trace(int device, const double* A, int64_t n, int ctas, long long* trace_out);

/* ---- horizon > 1 (dynamic GPIRT), constant_IRF == 0 ----
 * A panel is `horizon` single-session samplers (given theta the sessions are independent: the reference loops h outside every per-item
 * block) plus the joint theta step of src/draw-theta.cpp:171-236. cfg->horizon = H; regime by cfg->theta_ls as in the reference:
 *   theta_ls <= 0.1           RDM: every (respondent, session) its own ESS, prior sd sqrt(1 + theta_prior_sds(0,i)^2)
 *   0.1 < theta_ls < 3H       GP over time: one H-dim ESS per respondent, nu = L_time z, L_time = chol(K_time(0..H-1) + 1e-6 I)
 *   theta_ls >= 3H, H > 1     refused (GPIRT_ERR_UNSUPPORTED): the reference reads thresholds out of bounds there and aborts
 * Session handles are borrowed with gpirt_panel_session and take every per-session call above (set_data / set_priors / set_state /
 * get_state / draw_f / draw_fstar / gather_f / draw_beta / draw_threshold / ll_bar ...) on slice h of the reference's cubes; do not
 * destroy them or call gpirt_draw_theta / gpirt_sweep on them. Philox streams are keyed by (chain, session). world == 1 only. */
typedef struct gpirt_panel gpirt_panel;
int gpirt_panel_create(const gpirt_config* cfg, gpirt_panel** out);
int gpirt_panel_destroy(gpirt_panel* p);
int gpirt_panel_session(gpirt_panel* p, int h, gpirt_sampler** out);
int gpirt_panel_get_L_time(gpirt_panel* p, double* out /* H x H */);
/* streams: NULL (Philox), or RDM: an array of H gpirt_streams (session h: 1 normal + uniforms per respondent); GP: one gpirt_streams
 * with z_stride >= H normals per respondent */
int gpirt_panel_draw_theta(gpirt_panel* p, const gpirt_streams* streams);
int gpirt_panel_sweep(gpirt_panel* p, int n_sweeps);
int gpirt_panel_total_ll(gpirt_panel* p, double* total);
int gpirt_panel_sync(gpirt_panel* p);

/* ---- the two Rcpp entry points ---- */
typedef struct {
    /* inputs mirror gpirtMCMC()'s arguments (src/gpirtMCMC.cpp:25-38) */
    const double* y; int32_t n, m, horizon;
    const double* theta;                       /* n x horizon initial values */
    int32_t sample_iterations, burn_iterations, THIN;
    const double* beta_prior_means;            /* 3 x m */
    const double* beta_prior_sds;              /* 3 x m */
    const double* theta_prior_means;           /* 2 x n (unused by the sampler, as in the reference) */
    const double* theta_prior_sds;             /* 2 x n */
    double theta_os, theta_ls; const char* KERNEL;
    const double* thresholds; int32_t C;       /* m x (C+1) x horizon */
    int32_t constant_IRF, store_f, store_fstar;
    /* the reference draws these with R's RNG inside the call (src/gpirtMCMC.cpp:149-163); the R shim draws
     * them and passes them in; NULL => drawn on the device from the Philox INIT streams */
    const double* beta_init;                   /* 3 x m or NULL */
    const double* f_init;                      /* n x m or NULL */
    uint64_t seed; uint32_t chain; int32_t device;
    /* outputs, caller-allocated (n_samples = sample_iterations / THIN) */
    double* theta_draws;                       /* n_samples x n x horizon */
    double* beta_draws;                        /* n_samples slices of 3 x m x horizon */
    double* threshold_draws;                   /* n_samples slices of m x (C+1) x horizon */
    double* f_draws;                           /* n_samples slices of n x m x horizon, or NULL */
    double* fstar_draws;                       /* n_samples slices of NGRID x m x horizon, or NULL */
    double* ll_draws;                          /* n_samples */
    /* progress / interrupt poll, called once per sweep on the calling thread (Rprintf + checkUserInterrupt,
     * src/gpirtMCMC.cpp:262-264); return non-zero to stop. May be NULL. */
    int (*progress)(int iter, int total, void* user); void* user;
    /* n x m N(0,1) draws or NULL: f0_j = L z_j is formed on the device (rmvnorm(L), src/mvnormal.h:7-14, src/gpirtMCMC.cpp:158-162), so
     * an R shim that draws z with R::rnorm starts from the reference's own initial state without factorising K on the host.
     * Ignored when f_init is given. */
    const double* f_init_z;
    /* Several GPUs behind the same call (the north star's "work is partitioned across the 8 x B200 box by items"): n_devices > 1 runs the chain
     * item-sharded over devices[0..n_devices-1] from this one process - one shard per entry (an ordinal may repeat), peer copies for the exchanges,
     * the item-sharded theta step - and returns the draws of the single-device chain (same Philox streams; decisions identical unless a margin is
     * below fp64 summation noise). 0 or 1: `device` alone, unless the environment variable GPIRT_DEVICES ("0,1,2,3") names several - the way an
     * unmodified caller (the R shim, src/RcppExports.cpp:16-39) reaches them. Needs horizon == 1, constant_IRF == 0, no f / f* draws kept. */
    int32_t n_devices; const int32_t* devices;
} gpirt_mcmc_args;
/* constant_IRF == 1 (any horizon; with one session the reference's branch is the ordinary sampler with the inducing-point f* draw): composed from a panel, a stacked sampler over the n*horizon rows and a 100-respondent inducing-point sampler
 * (csrc/entry.cu mcmc_constant; src/draw-f.cpp:84-139, src/draw-fstar.cpp:58-125, src/draw_threshold.cpp:5-96). beta_init / f_init / f_init_z are
 * then read for session 0 only (the reference's constant branch draws nothing else, src/gpirtMCMC.cpp:176-188). */
int gpirt_mcmc(const gpirt_mcmc_args* a);
/* The memory estimate gpirtMCMC() prints before it allocates the kept draws (src/gpirtMCMC.cpp:47-82): MB of the theta / beta / f / f* / threshold
 * draws for n_samples = sample_iterations / THIN, their total over what store_f / store_fstar keep, and the reference's warning when the total
 * exceeds 10 000 MB. Host arithmetic only (no device needed). `text` (may be NULL) receives the block the reference writes to Rcpp::Rcout, same
 * wording and number format; the R shim prints it there. warn != 0: the total exceeds 10 000 MB. */
typedef struct { int32_t n_samples; double theta_mb, beta_mb, f_mb, fstar_mb, threshold_mb, total_mb; int32_t warn; } gpirt_mem_estimate;
int gpirt_memory_estimate(const gpirt_mcmc_args* a, gpirt_mem_estimate* out, char* text, int64_t text_cap);
/* The chain loop of the R wrapper (R/gpirtMCMC.R:111-158: `for(chain in 1:CHAIN)`, one .gpirtMCMC call per chain, run back to back) as ONE
 * call that runs the chains side by side on the device - BASELINE config 5 "independent chains batched per GPU". chains[c] is chain c's
 * argument block: same y (the first chain's is uploaded once and shared on the device), n, m, horizon == 1, C, iteration counts, THIN, KERNEL
 * and device; its own theta / beta_init / f_init / f_init_z, priors, seed, chain id and output buffers. Sweeps are issued round-robin, every
 * chain on its own pair of streams, so one chain's latency-bound Cholesky chain runs under the others' throughput-bound item kernels. Each
 * chain's draws are bit-identical to gpirt_mcmc() on the same argument block (Philox streams are keyed by chain id, kernels are
 * deterministic). progress is polled from chains[0] once per sweep. */
int gpirt_mcmc_chains(const gpirt_mcmc_args* chains, int32_t n_chains);

typedef struct {
    int32_t seed; const double* f; const double* y; int32_t n, m, horizon;
    const double* theta; const double* beta /* 2 x m x horizon */; const double* thresholds; int32_t C;
    const double* beta_prior_means; const double* beta_prior_sds; int32_t constant_IRF; int32_t device;
    double* fstar /* NGRID x m x horizon */;
} gpirt_recover_fstar_args;
/* One draw_f + draw_fstar from a stored f, RNG seeded with `seed` (src/recover_fstar.cpp:80-90). f, y: n x m x horizon; theta: n x horizon;
 * thresholds: m x (C+1) x horizon; beta_prior_sds: 2 x m. constant_IRF == 0: the sessions are drawn one after the other, session h on the
 * Philox streams of chain id h. constant_IRF == 1: the stacked draw_f with mu_h = [1, theta_h] beta_h, f interpolated onto 100
 * inducing points, one f* copied to every slice (src/draw-f.cpp:84-139, src/draw-fstar.cpp:58-125). */
int gpirt_recover_fstar(const gpirt_recover_fstar_args* a);

#ifdef __cplusplus
}
#endif
#endif
