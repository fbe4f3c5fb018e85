// The two Rcpp entry points of the reference as C-ABI calls with host buffers:
//   gpirt_mcmc            <- gpirtMCMC()      src/gpirtMCMC.cpp:25-406
//   gpirt_recover_fstar   <- recover_fstar()  src/recover_fstar.cpp:8-93
// Host side only: sample storage, thinning and the progress/interrupt poll; all arithmetic runs in the
// kernels behind the subsystem calls.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <algorithm>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../../include/gpirt_b200.h"
namespace gpirt { int fail_msg(int code, const char* msg); }  // capi.cu: sets gpirt_last_error()

extern "C" int gpirt_init_draw(gpirt_sampler* h, const double* beta_prior_means, int draw_beta, int draw_f);

namespace {
struct Guard {
    gpirt_sampler* h = nullptr;
    ~Guard() { if (h) gpirt_destroy(h); }
};
}  // namespace

#define TRYC(x)                 \
    do {                        \
        int rc_ = (x);          \
        if (rc_ != GPIRT_OK) return rc_; \
    } while (0)


// horizon > 1: the same run over a gpirt_panel. Every per-session array is slice h of the reference's cubes (column-major, slices contiguous).
static int mcmc_panel(const gpirt_mcmc_args* a, const gpirt_config& cfg) {
    struct PanelGuard { gpirt_panel* p = nullptr; ~PanelGuard() { if (p) gpirt_panel_destroy(p); } } g;
    TRYC(gpirt_panel_create(&cfg, &g.p));
    const int H = a->horizon;
    const size_t n = a->n, m = a->m, C1 = (size_t)a->C + 1;
    std::vector<gpirt_sampler*> S(H);
    for (int h = 0; h < H; ++h) {
        TRYC(gpirt_panel_session(g.p, h, &S[h]));
        TRYC(gpirt_set_data_f64(S[h], a->y + (size_t)h * n * m));
        TRYC(gpirt_set_priors(S[h], a->beta_prior_sds, a->theta_prior_sds));
        TRYC(gpirt_set_state(S[h], a->theta + (size_t)h * n, a->beta_init ? a->beta_init + (size_t)h * 3 * m : nullptr,
                             a->f_init ? a->f_init + (size_t)h * n * m : nullptr, a->thresholds + (size_t)h * m * C1, nullptr));
        TRYC(gpirt_set_iter(S[h], 0xFFFFFFFFu));
        TRYC(gpirt_init_draw(S[h], a->beta_prior_means, a->beta_init == nullptr, a->f_init == nullptr && a->f_init_z == nullptr));
        if (!a->f_init && a->f_init_z) TRYC(gpirt_init_f_from_z(S[h], a->f_init_z + (size_t)h * n * m));
        TRYC(gpirt_set_iter(S[h], 0));
    }
    const int total = a->sample_iterations + a->burn_iterations;
    const int n_samples = a->sample_iterations / a->THIN;
    int cap = 0, first = 0;
    if (n_samples > 0) {
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return GPIRT_ERR_CUDA;
        const int64_t per = gpirt_store_sample_bytes(S[0], a->f_draws != nullptr, a->fstar_draws != nullptr) * H;
        cap = (int)std::max<int64_t>(1, std::min<int64_t>(n_samples, (int64_t)(free_b / 4) / per));
        for (int h = 0; h < H; ++h) TRYC(gpirt_store_open(S[h], cap, a->f_draws != nullptr, a->fstar_draws != nullptr));
    }
    auto flush = [&]() -> int {
        int cnt = 0;
        TRYC(gpirt_store_count(S[0], &cnt, nullptr));
        for (int h = 0; h < H; ++h)
            TRYC(gpirt_store_flush_slice(S[h], n_samples, first, H, h, a->theta_draws, a->beta_draws, a->threshold_draws, a->ll_draws, a->f_draws,
                                         a->fstar_draws));
        first += cnt;
        return GPIRT_OK;
    };
    for (int iter = 0; iter < total; ++iter) {
        if (a->progress && a->progress(iter, total, a->user)) return GPIRT_ERR_INVALID;
        TRYC(gpirt_panel_sweep(g.p, 1));
        if (iter >= a->burn_iterations && iter % a->THIN == 0) {
            const int s = (iter - a->burn_iterations) / a->THIN;
            if (s >= n_samples) continue;
            int cnt = 0;
            for (int h = 0; h < H; ++h) TRYC(gpirt_store_push(S[h]));
            TRYC(gpirt_store_count(S[0], &cnt, nullptr));
            if (cnt == cap) TRYC(flush());
        }
    }
    if (n_samples > 0) TRYC(flush());
    return gpirt_panel_sync(g.p);
}

// constant_IRF == 1 (src/draw-f.cpp:84-139, src/draw-fstar.cpp:58-125, src/draw_threshold.cpp:5-96,181-205, init src/gpirtMCMC.cpp:164-200) as a
// composition of the handles above: the panel's H sessions (theta step, gather, draw_beta, log-likelihood — no constant_IRF argument in the
// reference), a STACKED sampler T over the n*H (session, respondent) rows whose ordinary draw_f / draw_threshold are the constant-IRF blocks once
// the rows carry their own session's mean (gpirt_set_mean_rows), and a 100-respondent sampler I whose ordinary draw_fstar is the inducing-point
// posterior once theta is the inducing grid and f the stacked f interpolated onto it (gpirt_interp1). State moves between the handles through
// host buffers: every likelihood / ESS / factorisation runs in the kernels, the mode is not tuned. Same composition as
// gpirt_b200/constant_panel.py (bit-identical chains, tests/test_gpu_panel.py).
static int mcmc_constant(const gpirt_mcmc_args* a, gpirt_config cfg) {
    const int H = a->horizon, P = 100;  // n_induced_points, src/draw-fstar.cpp:73
    const size_t n = a->n, m = a->m, C1 = (size_t)a->C + 1, N = 1001;
    // H == 1 is a legal call in the reference too: the ordinary sampler with the f* draw taken through the inducing points
    cfg.constant_IRF = 0;  // the handles are the ordinary ones
    struct Owner {
        gpirt_panel* p = nullptr; gpirt_sampler* T = nullptr; gpirt_sampler* I = nullptr;
        ~Owner() { if (T) gpirt_destroy(T); if (I) gpirt_destroy(I); if (p) gpirt_panel_destroy(p); }
    } o;
    TRYC(gpirt_panel_create(&cfg, &o.p));
    std::vector<gpirt_sampler*> S(H);
    for (int h = 0; h < H; ++h) TRYC(gpirt_panel_session(o.p, h, &S[h]));
    gpirt_config ct = cfg; ct.horizon = 1; ct.n = (int32_t)(n * H);
    TRYC(gpirt_create(&ct, &o.T));
    gpirt_config ci = cfg; ci.horizon = 1; ci.n = P;
    TRYC(gpirt_create(&ci, &o.I));
    // data and priors
    const double nan = std::numeric_limits<double>::quiet_NaN();
    std::vector<double> ystack(n * H * m), tps_stack(2 * n * H), zeros_i(2 * (size_t)P, 0.0), y_i((size_t)P * m, nan);
    for (size_t j = 0; j < m; ++j)
        for (int h = 0; h < H; ++h) std::memcpy(&ystack[j * n * H + h * n], a->y + (size_t)h * n * m + j * n, n * 8);  // row h*n + i, src/draw-f.cpp:94-103
    for (int h = 0; h < H; ++h) std::memcpy(&tps_stack[(size_t)h * 2 * n], a->theta_prior_sds, 2 * n * 8);
    for (int h = 0; h < H; ++h) {
        TRYC(gpirt_set_data_f64(S[h], a->y + (size_t)h * n * m));
        TRYC(gpirt_set_priors(S[h], a->beta_prior_sds, a->theta_prior_sds));
    }
    TRYC(gpirt_set_data_f64(o.T, ystack.data()));
    TRYC(gpirt_set_priors(o.T, a->beta_prior_sds, tps_stack.data()));
    TRYC(gpirt_set_data_f64(o.I, y_i.data()));
    TRYC(gpirt_set_priors(o.I, a->beta_prior_sds, zeros_i.data()));
    // initialisation, constant branch: beta for session 0 only, one f0 = L_0 z per item for every session, every session's mu = X_h beta_0
    std::vector<double> th(n * H), beta0(3 * m), f0(n * m), zero_beta(3 * m, 0.0), mu(n * m * H);
    for (size_t k = 0; k < n * H; ++k) th[k] = std::min(5.0, std::max(-5.0, a->theta[k]));  // src/gpirtMCMC.cpp:133
    auto set_iter_all = [&](uint32_t it) -> int {
        for (int h = 0; h < H; ++h) TRYC(gpirt_set_iter(S[h], it));
        TRYC(gpirt_set_iter(o.T, it));
        return gpirt_set_iter(o.I, it);
    };
    TRYC(gpirt_set_state(S[0], th.data(), a->beta_init, a->f_init, a->thresholds, nullptr));
    TRYC(gpirt_set_iter(S[0], 0xFFFFFFFFu));
    TRYC(gpirt_init_draw(S[0], a->beta_prior_means, a->beta_init == nullptr, a->f_init == nullptr && a->f_init_z == nullptr));
    if (!a->f_init && a->f_init_z) TRYC(gpirt_init_f_from_z(S[0], a->f_init_z));
    TRYC(gpirt_get_state(S[0], nullptr, beta0.data(), f0.data(), nullptr, nullptr));
    auto x_beta = [&](const double* theta, const double* beta, double* out) {  // mu = X beta, left to right (src/gpirtMCMC.cpp:308)
        for (size_t j = 0; j < m; ++j)
            for (size_t i = 0; i < n; ++i) {
                const double t = theta[i];
                out[j * n + i] = (beta[3 * j] + t * beta[3 * j + 1]) + (t * t) * beta[3 * j + 2];
            }
    };
    for (int h = 0; h < H; ++h) {
        TRYC(gpirt_set_state(S[h], &th[(size_t)h * n], h == 0 ? beta0.data() : zero_beta.data(), f0.data(), a->thresholds + (size_t)h * m * C1, nullptr));
        x_beta(&th[(size_t)h * n], beta0.data(), &mu[(size_t)h * n * m]);
    }
    // stacked views of the sessions' state
    std::vector<double> th_s(n * H), f_s(n * H * m), mu_s(n * H * m), f_h(n * m), beta_h(3 * m), thr0(m * C1), fs(N * m), tind(P), find((size_t)P * m);
    auto stack = [&]() -> int {
        for (int h = 0; h < H; ++h) {
            TRYC(gpirt_get_state(S[h], &th_s[(size_t)h * n], nullptr, f_h.data(), nullptr, nullptr));
            for (size_t j = 0; j < m; ++j) {
                std::memcpy(&f_s[j * n * H + h * n], &f_h[j * n], n * 8);
                std::memcpy(&mu_s[j * n * H + h * n], &mu[(size_t)h * n * m + j * n], n * 8);
            }
        }
        TRYC(gpirt_get_state(S[0], nullptr, beta_h.data(), nullptr, thr0.data(), nullptr));
        return GPIRT_OK;
    };
    auto draw_f = [&]() -> int {
        TRYC(stack());
        TRYC(gpirt_set_state(o.T, th_s.data(), beta_h.data(), f_s.data(), thr0.data(), nullptr));
        TRYC(gpirt_set_mean_rows(o.T, mu_s.data()));
        TRYC(gpirt_draw_f(o.T, nullptr));
        TRYC(gpirt_get_state(o.T, nullptr, nullptr, f_s.data(), nullptr, nullptr));
        for (int h = 0; h < H; ++h) {
            for (size_t j = 0; j < m; ++j) std::memcpy(&f_h[j * n], &f_s[j * n * H + h * n], n * 8);
            TRYC(gpirt_set_state(S[h], nullptr, nullptr, f_h.data(), nullptr, nullptr));
        }
        return GPIRT_OK;
    };
    auto draw_fstar = [&]() -> int {
        TRYC(stack());
        const double lo = *std::min_element(th_s.begin(), th_s.end()), hi = *std::max_element(th_s.begin(), th_s.end());
        const double delta = (hi - lo) / double(P - 1);  // arma::linspace: start + i * delta, last element = end
        for (int i = 0; i + 1 < P; ++i) tind[i] = lo + double(i) * delta;
        tind[P - 1] = hi;
        TRYC(gpirt_interp1(cfg.device, th_s.data(), (int64_t)(n * H), f_s.data(), (int64_t)m, tind.data(), P, find.data()));  // src/draw-fstar.cpp:77-82
        TRYC(gpirt_set_state(o.I, tind.data(), beta_h.data(), find.data(), nullptr, nullptr));
        TRYC(gpirt_draw_fstar(o.I, nullptr));
        TRYC(gpirt_get_state(o.I, nullptr, nullptr, nullptr, nullptr, fs.data()));
        for (int h = 0; h < H; ++h) TRYC(gpirt_set_state(S[h], nullptr, nullptr, nullptr, nullptr, fs.data()));  // one f* for all sessions (:121-124)
        return GPIRT_OK;
    };
    auto draw_threshold = [&]() -> int {
        TRYC(stack());
        TRYC(gpirt_set_state(o.T, th_s.data(), nullptr, f_s.data(), thr0.data(), nullptr));
        TRYC(gpirt_set_mean_rows(o.T, mu_s.data()));
        TRYC(gpirt_draw_threshold(o.T, nullptr));
        TRYC(gpirt_get_state(o.T, nullptr, nullptr, nullptr, thr0.data(), nullptr));
        for (int h = 0; h < H; ++h) TRYC(gpirt_set_state(S[h], nullptr, nullptr, nullptr, thr0.data(), nullptr));  // src/draw_threshold.cpp:199-203
        return GPIRT_OK;
    };
    TRYC(set_iter_all(0xFFFFFFFFu));
    TRYC(draw_fstar());  // src/gpirtMCMC.cpp:227
    TRYC(set_iter_all(0));
    const int total = a->sample_iterations + a->burn_iterations;
    const int n_samples = a->sample_iterations / a->THIN;
    for (int iter = 0; iter < total; ++iter) {
        if (a->progress && a->progress(iter, total, a->user)) return GPIRT_ERR_INVALID;
        TRYC(draw_f());
        TRYC(draw_fstar());
        TRYC(gpirt_panel_draw_theta(o.p, nullptr));
        for (int h = 0; h < H; ++h) TRYC(gpirt_gather_f(S[h]));
        for (int h = 0; h < H; ++h) TRYC(gpirt_draw_beta(S[h], nullptr));
        for (int h = 0; h < H; ++h) {  // mu.slice(h) = X.slice(h) * beta.slice(h), src/gpirtMCMC.cpp:307-311
            TRYC(gpirt_get_state(S[h], &th_s[(size_t)h * n], beta_h.data(), nullptr, nullptr, nullptr));
            x_beta(&th_s[(size_t)h * n], beta_h.data(), &mu[(size_t)h * n * m]);
            TRYC(gpirt_update_cholesky_cache(S[h]));
        }
        TRYC(draw_threshold());
        TRYC(set_iter_all((uint32_t)iter + 1));
        if (iter >= a->burn_iterations && iter % a->THIN == 0) {  // src/gpirtMCMC.cpp:334-335
            const size_t s = (size_t)(iter - a->burn_iterations) / a->THIN;
            if ((int)s >= n_samples) continue;
            double ll = 0.0;
            for (int h = 0; h < H; ++h) {
                double part = 0.0;
                TRYC(gpirt_ll_bar(S[h], nullptr, &part));
                ll += part;
                TRYC(gpirt_get_state(S[h], &th_s[(size_t)h * n], a->beta_draws + s * 3 * m * H + (size_t)h * 3 * m, a->f_draws ? a->f_draws + s * n * m * H + (size_t)h * n * m : nullptr,
                                     a->threshold_draws + s * m * C1 * H + (size_t)h * m * C1, a->fstar_draws ? a->fstar_draws + s * N * m * H + (size_t)h * N * m : nullptr));
                for (size_t i = 0; i < n; ++i) a->theta_draws[s + (size_t)n_samples * (i + n * h)] = th_s[(size_t)h * n + i];
            }
            a->ll_draws[s] = ll;
        }
    }
    return gpirt_panel_sync(o.p);
}

static bool args_ok(const gpirt_mcmc_args* a) {
    if (!a || !a->y || !a->theta || !a->beta_prior_sds || !a->theta_prior_sds || !a->thresholds || !a->theta_draws || !a->beta_draws ||
        !a->threshold_draws || !a->ll_draws)
        return false;
    return a->THIN >= 1 && a->sample_iterations >= 0 && a->burn_iterations >= 0;
}

static gpirt_config config_of(const gpirt_mcmc_args* a) {
    gpirt_config cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.n = a->n; cfg.m_total = a->m; cfg.horizon = a->horizon; cfg.C = a->C; cfg.constant_IRF = a->constant_IRF; cfg.p_beta = 3;
    cfg.theta_os = a->theta_os; cfg.theta_ls = a->theta_ls;
    const std::string kern = a->KERNEL ? a->KERNEL : "Matern";
    cfg.kernel = kern == "RBF" ? 1 : 0;
    cfg.device = a->device; cfg.seed = a->seed; cfg.chain = a->chain; cfg.rank = 0; cfg.world = 1;
    return cfg;
}

namespace {
// One chain of gpirtMCMC() at horizon 1: the state before the loop (src/gpirtMCMC.cpp:133-227), one sweep + store decision per step()
// (:261-356), bulk read-back of the kept draws. Nothing in step() synchronises with the host, so several chains can be stepped round-robin
// from one thread and overlap on the device.
struct ChainRun {
    const gpirt_mcmc_args* a = nullptr;
    Guard g;
    int n_samples = 0, cap = 0, first = 0;

    int open(const gpirt_mcmc_args* args, gpirt_sampler* data_from, int n_chains) {
        a = args;
        gpirt_config cfg = config_of(a);
        TRYC(gpirt_create(&cfg, &g.h));
        gpirt_sampler* h = g.h;
        TRYC(gpirt_set_priors(h, a->beta_prior_sds, a->theta_prior_sds));
        TRYC(gpirt_set_state(h, a->theta, a->beta_init, a->f_init, a->thresholds, nullptr));
        TRYC(gpirt_set_iter(h, 0xFFFFFFFFu));  // initial draws use a sweep id no sweep reaches
        TRYC(gpirt_init_draw(h, a->beta_prior_means, a->beta_init == nullptr, a->f_init == nullptr && a->f_init_z == nullptr));
        if (!a->f_init && a->f_init_z) TRYC(gpirt_init_f_from_z(h, a->f_init_z));
        TRYC(gpirt_set_iter(h, 0));
        // the responses last: the initial draws (K, its Cholesky, f0 = L z) do not read them and are already running when their 8 n m bytes start
        // over PCIe on the handle's side stream (gpirt_set_data_f64)
        if (data_from) TRYC(gpirt_copy_data(h, data_from));
        else TRYC(gpirt_set_data_f64(h, a->y));
        n_samples = a->sample_iterations / a->THIN;  // src/gpirtMCMC.cpp:48
        if (n_chains > 1) {  // chains side by side share the SMs: the persistent Cholesky's CTA budget is divided among them. The factor does not depend
            // on the CTA count (updates are applied in step order), so every chain stays bit-identical to the chain run alone; the per-panel path would not
            static const int budget = getenv("GPIRT_CHAINS_CHOL_CTAS") ? atoi(getenv("GPIRT_CHAINS_CHOL_CTAS")) : 96;
            TRYC(gpirt_set_chol_ctas(h, std::max(6, budget / n_chains)));
        }
        return GPIRT_OK;
    }
    // The kept draws stay on the device and come back in bulk: no host synchronisation per sweep, so the next sweep's theta-only
    // chain keeps running under this sweep's tail. Capacity: as many samples as fit in this chain's share of a quarter of the free memory.
    int open_store(size_t budget_bytes) {
        if (n_samples <= 0) return GPIRT_OK;
        const int64_t per = gpirt_store_sample_bytes(g.h, a->f_draws != nullptr, a->fstar_draws != nullptr);
        cap = (int)std::max<int64_t>(1, std::min<int64_t>(n_samples, (int64_t)budget_bytes / per));
        return gpirt_store_open(g.h, cap, a->f_draws != nullptr, a->fstar_draws != nullptr);
    }
    int flush() {
        int cnt = 0;
        TRYC(gpirt_store_count(g.h, &cnt, nullptr));
        TRYC(gpirt_store_flush(g.h, n_samples, first, a->theta_draws, a->beta_draws, a->threshold_draws, a->ll_draws, a->f_draws, a->fstar_draws));
        first += cnt;
        return GPIRT_OK;
    }
    int step(int iter) {
        TRYC(gpirt_sweep(g.h, 1));
        if (iter >= a->burn_iterations && iter % a->THIN == 0) {  // src/gpirtMCMC.cpp:334-335
            const int s = (iter - a->burn_iterations) / a->THIN;
            if (s >= n_samples) return GPIRT_OK;  // the reference would write past theta_draws here
            TRYC(gpirt_store_push(g.h));
            int cnt = 0;
            TRYC(gpirt_store_count(g.h, &cnt, nullptr));
            if (cnt == cap) TRYC(flush());
        }
        return GPIRT_OK;
    }
    int close() {
        if (n_samples > 0) TRYC(flush());
        return gpirt_sync(g.h);
    }
};
}  // namespace

// gpirtMCMC() on several GPUs from ONE process (gpirt_mcmc_args.n_devices / GPIRT_DEVICES): the item-sharded chain of gpirt_b200/distributed.py with
// the exchanges as peer copies (gpirt_exchange_pull). Shard r = handle of rank r on devs[r]. Per sweep (src/gpirtMCMC.cpp:267-331):
//   draw_f, draw_fstar on every shard's items | theta rounds { partials ; pull the other shards' partials ; decide } | f <- f*, draw_beta, K +
//   Cholesky (replicated), draw_threshold, item ll on every shard | one packed exchange of beta / thresholds / item ll.
// The kept draws (theta, beta, thresholds, ll: complete on every shard after the packed exchange) are stored on shard 0.
static int mcmc_sharded(const gpirt_mcmc_args* a, const std::vector<int>& devs) {
    const int W = (int)devs.size();
    if (a->horizon != 1 || a->constant_IRF != 0) return gpirt::fail_msg(GPIRT_ERR_UNSUPPORTED, "several devices: horizon 1 and constant_IRF 0 only");
    if (a->f_draws || a->fstar_draws) return gpirt::fail_msg(GPIRT_ERR_UNSUPPORTED, "several devices: f / f* draws are not kept (every shard holds its own items only)");
    if (W > a->m || W > a->n) return gpirt::fail_msg(GPIRT_ERR_SHAPE, "several devices: more shards than items or respondents");
    for (int r = 0; r < W; ++r)  // direct NVLink copies between the shards' GPUs (already enabled / same device: fine)
        for (int q = 0; q < W; ++q)
            if (devs[r] != devs[q]) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, devs[r], devs[q]) == cudaSuccess && can && cudaSetDevice(devs[r]) == cudaSuccess) {
                    cudaDeviceEnablePeerAccess(devs[q], 0);
                    cudaGetLastError();
                }
            }
    std::vector<Guard> g(W);
    std::vector<gpirt_sampler*> h(W, nullptr);
    for (int r = 0; r < W; ++r) {
        gpirt_config cfg = config_of(a);
        cfg.device = devs[r]; cfg.rank = r; cfg.world = W;
        TRYC(gpirt_create(&cfg, &g[r].h));
        h[r] = g[r].h;
        TRYC(gpirt_set_priors(h[r], a->beta_prior_sds, a->theta_prior_sds));
        TRYC(gpirt_set_state(h[r], a->theta, a->beta_init, a->f_init, a->thresholds, nullptr));
        TRYC(gpirt_set_iter(h[r], 0xFFFFFFFFu));
        TRYC(gpirt_init_draw(h[r], a->beta_prior_means, a->beta_init == nullptr, a->f_init == nullptr && a->f_init_z == nullptr));
        if (!a->f_init && a->f_init_z) TRYC(gpirt_init_f_from_z(h[r], a->f_init_z));
        TRYC(gpirt_set_iter(h[r], 0));
        TRYC(gpirt_set_data_f64(h[r], a->y));  // a shard uploads its items' columns and its respondents' rows, under its initial draws
    }
    {   // config-4 sizes with every shard on a GPU of its own: the theta-only chain is distributed over the shards (gpirt_chain_*, same process: no IPC)
        bool distinct = true;
        for (int r = 0; r < W; ++r)
            for (int q = 0; q < r; ++q) distinct = distinct && devs[r] != devs[q];
        if (distinct && a->n >= 4096) {
            std::vector<int64_t> offs((size_t)W * 6);
            unsigned char dummy[64];
            for (int r = 0; r < W; ++r) TRYC(gpirt_chain_export(h[r], dummy, &offs[(size_t)r * 6]));
            for (int r = 0; r < W; ++r)
                for (int q = 0; q < W; ++q)
                    if (q != r) TRYC(gpirt_chain_attach(h[r], q, nullptr, &offs[(size_t)q * 6]));
        }
    }
    auto exchange = [&](int which) -> int {
        for (int q = 0; q < W; ++q)
            for (int r = 0; r < W; ++r)
                if (r != q) TRYC(gpirt_exchange_pull(h[q], h[r], which));
        return GPIRT_OK;
    };
    const int n_samples = a->sample_iterations / a->THIN;  // src/gpirtMCMC.cpp:48
    int cap = 0, first = 0;
    if (n_samples > 0) {
        size_t free_b = 0, total_b = 0;
        if (cudaSetDevice(devs[0]) != cudaSuccess || cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return GPIRT_ERR_CUDA;
        const int64_t per = gpirt_store_sample_bytes(h[0], 0, 0);
        cap = (int)std::max<int64_t>(1, std::min<int64_t>(n_samples, (int64_t)(free_b / 4) / per));
        TRYC(gpirt_store_open(h[0], cap, 0, 0));
    }
    auto flush = [&]() -> int {
        int cnt = 0;
        TRYC(gpirt_store_count(h[0], &cnt, nullptr));
        TRYC(gpirt_store_flush(h[0], n_samples, first, a->theta_draws, a->beta_draws, a->threshold_draws, a->ll_draws, nullptr, nullptr));
        first += cnt;
        return GPIRT_OK;
    };
    const int total = a->sample_iterations + a->burn_iterations;
    std::vector<int32_t> active(W, 0);
    for (int iter = 0; iter < total; ++iter) {
        if (a->progress && a->progress(iter, total, a->user)) return GPIRT_ERR_INVALID;  // user interrupt
        for (int r = 0; r < W; ++r) TRYC(gpirt_sweep_phase_items(h[r]));
        for (int r = 0; r < W; ++r) TRYC(gpirt_theta_items_begin(h[r], nullptr));
        for (;;) {
            for (int r = 0; r < W; ++r) TRYC(gpirt_theta_items_partials(h[r]));
            TRYC(exchange(GPIRT_XCHG_THETA_PART));
            for (int r = 0; r < W; ++r) TRYC(gpirt_theta_items_decide(h[r], &active[r]));
            for (int r = 1; r < W; ++r)
                if (active[r] != active[0]) return gpirt::fail_msg(GPIRT_ERR_CUDA, "several devices: the shards disagree on the unfinished respondents");
            if (active[0] == 0) break;
        }
        for (int r = 0; r < W; ++r) TRYC(gpirt_sweep_phase_close(h[r]));
        for (int r = 0; r < W; ++r) TRYC(gpirt_items_pack(h[r], 0));
        TRYC(exchange(GPIRT_XCHG_ITEMS_PACKED));
        for (int r = 0; r < W; ++r) TRYC(gpirt_items_pack(h[r], 1));
        if (iter >= a->burn_iterations && iter % a->THIN == 0) {  // src/gpirtMCMC.cpp:334-335
            const int sidx = (iter - a->burn_iterations) / a->THIN;
            if (sidx < n_samples) {
                TRYC(gpirt_store_push(h[0]));
                int cnt = 0;
                TRYC(gpirt_store_count(h[0], &cnt, nullptr));
                if (cnt == cap) TRYC(flush());
            }
        }
    }
    if (n_samples > 0) TRYC(flush());
    for (int r = 0; r < W; ++r) TRYC(gpirt_sync(h[r]));
    for (int r = 0; r < W; ++r) TRYC(gpirt_chain_detach(h[r]));  // before any shard's arena is freed
    return GPIRT_OK;
}

// the devices of a call: args.devices, else GPIRT_DEVICES ("0,1,2,3"), else args.device alone
static std::vector<int> devices_of(const gpirt_mcmc_args* a) {
    std::vector<int> d;
    if (a->n_devices > 1 && a->devices) {
        d.assign(a->devices, a->devices + a->n_devices);
    } else if (a->n_devices <= 0) {
        if (const char* e = getenv("GPIRT_DEVICES")) {
            for (const char* p = e; *p;) {
                char* end = nullptr;
                const long v = strtol(p, &end, 10);
                if (end == p) break;
                d.push_back((int)v);
                p = (*end == ',') ? end + 1 : end;
            }
        }
    }
    if (d.size() < 2) d.assign(1, a->device);
    return d;
}

extern "C" int gpirt_mcmc_chains(const gpirt_mcmc_args* chains, int32_t n_chains) {
    if (!chains || n_chains < 1) return GPIRT_ERR_INVALID;
    const gpirt_mcmc_args* a0 = &chains[0];
    for (int c = 0; c < n_chains; ++c) {
        const gpirt_mcmc_args* a = &chains[c];
        if (!args_ok(a)) return GPIRT_ERR_INVALID;
        if (a->n != a0->n || a->m != a0->m || a->horizon != a0->horizon || a->C != a0->C || a->device != a0->device ||
            a->sample_iterations != a0->sample_iterations || a->burn_iterations != a0->burn_iterations || a->THIN != a0->THIN)
            return GPIRT_ERR_SHAPE;
    }
    if (a0->horizon == 1 && !a0->constant_IRF) {  // several GPUs: every chain item-sharded over them, chains one after the other
        const std::vector<int> devs = devices_of(a0);
        if (devs.size() > 1) {
            for (int c = 0; c < n_chains; ++c) TRYC(mcmc_sharded(&chains[c], devs));
            return GPIRT_OK;
        }
    }
    if (a0->constant_IRF) {  // one item response function for all sessions: the composed mode, chains one after the other
        for (int c = 0; c < n_chains; ++c) TRYC(mcmc_constant(&chains[c], config_of(&chains[c])));
        return GPIRT_OK;
    }
    if (a0->horizon > 1) {  // dynamic GPIRT: chains one after the other (a panel already keeps H samplers in flight)
        for (int c = 0; c < n_chains; ++c) TRYC(mcmc_panel(&chains[c], config_of(&chains[c])));
        return GPIRT_OK;
    }
    std::vector<ChainRun> runs(n_chains);
    for (int c = 0; c < n_chains; ++c) {
        const bool same_y = c > 0 && chains[c].y == a0->y;
        TRYC(runs[c].open(&chains[c], same_y ? runs[0].g.h : nullptr, n_chains));
    }
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return GPIRT_ERR_CUDA;
    for (int c = 0; c < n_chains; ++c) TRYC(runs[c].open_store(free_b / 4 / (size_t)n_chains));
    const int total = a0->sample_iterations + a0->burn_iterations;
    for (int iter = 0; iter < total; ++iter) {
        if (a0->progress && a0->progress(iter, total, a0->user)) return GPIRT_ERR_INVALID;  // user interrupt
        for (int c = 0; c < n_chains; ++c) TRYC(runs[c].step(iter));
    }
    for (int c = 0; c < n_chains; ++c) TRYC(runs[c].close());
    return GPIRT_OK;
}

extern "C" int gpirt_memory_estimate(const gpirt_mcmc_args* a, gpirt_mem_estimate* out, char* text, int64_t text_cap) {
    if (!a || !out || a->THIN < 1 || a->n < 1 || a->m < 1 || a->horizon < 1) return gpirt::fail_msg(GPIRT_ERR_INVALID, "memory_estimate: bad arguments");
    // src/gpirtMCMC.cpp:48-58; its `C` is thresholds.n_cols, i.e. categories + 1
    const double n = a->n, m = a->m, H = a->horizon, C1 = a->C + 1, mb = 1024.0 * 1024.0;
    const int n_samples = a->sample_iterations / a->THIN;
    gpirt_mem_estimate e;
    e.n_samples = n_samples;
    e.theta_mb = (n_samples * n * H * 8.0) / mb;
    e.beta_mb = (n_samples * 3 * m * H * 8.0) / mb;
    e.f_mb = (n_samples * n * m * H * 8.0) / mb;
    e.fstar_mb = (n_samples * 1001 * m * H * 8.0) / mb;
    e.threshold_mb = (n_samples * m * C1 * H * 8.0) / mb;
    e.total_mb = e.theta_mb + e.beta_mb + e.threshold_mb;
    if (a->store_f) e.total_mb += e.f_mb;
    if (a->store_fstar) e.total_mb += e.fstar_mb;
    e.warn = e.total_mb > 10000 ? 1 : 0;
    *out = e;
    if (text && text_cap > 0) {  // :60-82, number format of an ostream with default precision ("%g")
        snprintf(text, (size_t)text_cap,
                 "\n=== MEMORY ESTIMATE ===\nSamples to store: %d (thinned from %d)\nTheta samples:     %g MB\nBeta samples:      %g MB\n"
                 "F samples:         %g MB (%s)\nFstar samples:     %g MB (%s)\nThreshold samples: %g MB\nTOTAL ESTIMATED:   %g MB (%g GB)\n%s"
                 "========================\n\n",
                 n_samples, a->sample_iterations, e.theta_mb, e.beta_mb, e.f_mb, a->store_f ? "ENABLED" : "DISABLED - will skip", e.fstar_mb,
                 a->store_fstar ? "ENABLED" : "DISABLED - will skip", e.threshold_mb, e.total_mb, e.total_mb / 1024.0,
                 e.warn ? "\nWARNING: Estimated memory usage exceeds 10 GB!\nConsider: (1) Increase THIN parameter, (2) Reduce sample_iterations\n"
                          "          (3) Set store_f=FALSE, (4) Set store_fstar=FALSE\n\n" : "");
    }
    return GPIRT_OK;
}

extern "C" int gpirt_mcmc(const gpirt_mcmc_args* a) {
    if (!args_ok(a)) return GPIRT_ERR_INVALID;
    return gpirt_mcmc_chains(a, 1);
}

// recover_fstar(): ws_pool.seed_all(seed), then draw_f and draw_fstar from the stored f (src/recover_fstar.cpp:80-90). Sessions are independent
// (constant_IRF == 0: one ordinary handle per session, Philox streams tagged with the session as the chain id) or share one item response
// function (constant_IRF == 1: the stacked sampler + inducing-point sampler of mcmc_constant, 2-row beta).
extern "C" int gpirt_recover_fstar(const gpirt_recover_fstar_args* a) {
    if (!a || !a->f || !a->y || !a->theta || !a->beta || !a->thresholds || !a->beta_prior_sds || !a->fstar) return GPIRT_ERR_INVALID;
    if (a->horizon < 1) return GPIRT_ERR_SHAPE;
    const int H = a->horizon;
    const size_t n = a->n, m = a->m, C1 = (size_t)a->C + 1, N = 1001;
    gpirt_config cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.n = a->n; cfg.m_total = a->m; cfg.horizon = 1; cfg.C = a->C; cfg.constant_IRF = 0; cfg.p_beta = 2;
    cfg.theta_os = 1.0; cfg.theta_ls = 10.0; cfg.kernel = 0; cfg.device = a->device;
    cfg.seed = (uint64_t)(uint32_t)a->seed;  // ws_pool.seed_all(seed), src/recover_fstar.cpp:80
    cfg.chain = 0; cfg.rank = 0; cfg.world = 1;
    // recover_fstar passes a 2 x m beta_prior_sds (R/recover_fstar.R:37-38); K() reads rows 0..2 of column 0
    // (src/covariance-function.cpp:12-14) — an out-of-bounds read in the reference. Row 2 is taken as 0 here.
    std::vector<double> sds3(3 * m, 0.0);
    for (size_t j = 0; j < m; ++j) { sds3[3 * j] = a->beta_prior_sds[2 * j]; sds3[3 * j + 1] = a->beta_prior_sds[2 * j + 1]; }
    if (!a->constant_IRF) {
        for (int h = 0; h < H; ++h) {
            Guard g;
            cfg.chain = (uint32_t)h;
            TRYC(gpirt_create(&cfg, &g.h));
            TRYC(gpirt_set_data_f64(g.h, a->y + (size_t)h * n * m));
            TRYC(gpirt_set_priors(g.h, sds3.data(), nullptr));
            TRYC(gpirt_set_state(g.h, a->theta + (size_t)h * n, a->beta + (size_t)h * 2 * m, a->f + (size_t)h * n * m, a->thresholds + (size_t)h * m * C1, nullptr));
            TRYC(gpirt_set_iter(g.h, 0));
            TRYC(gpirt_draw_f(g.h, nullptr));      // src/recover_fstar.cpp:87-88
            TRYC(gpirt_draw_fstar(g.h, nullptr));  // :89-90
            TRYC(gpirt_get_state(g.h, nullptr, nullptr, nullptr, nullptr, a->fstar + (size_t)h * N * m));
        }
        return GPIRT_OK;
    }
    // constant_IRF == 1: stacked draw_f (src/draw-f.cpp:84-139) with mu_h = [1, theta_h] beta_h, then the inducing-point f* draw (src/draw-fstar.cpp:58-125)
    const int P = 100;
    Guard T, I;
    gpirt_config ct = cfg; ct.n = (int32_t)(n * H);
    gpirt_config ci = cfg; ci.n = P;
    TRYC(gpirt_create(&ct, &T.h));
    TRYC(gpirt_create(&ci, &I.h));
    const double nan = std::numeric_limits<double>::quiet_NaN();
    std::vector<double> ys(n * H * m), fs(n * H * m), mus(n * H * m), ths(n * H), y_i((size_t)P * m, nan), tind(P), find((size_t)P * m), fstar(N * m);
    for (int h = 0; h < H; ++h)
        for (size_t i = 0; i < n; ++i) ths[(size_t)h * n + i] = a->theta[(size_t)h * n + i];
    for (size_t j = 0; j < m; ++j)
        for (int h = 0; h < H; ++h) {
            std::memcpy(&ys[j * n * H + h * n], a->y + (size_t)h * n * m + j * n, n * 8);
            std::memcpy(&fs[j * n * H + h * n], a->f + (size_t)h * n * m + j * n, n * 8);
            const double* b = a->beta + (size_t)h * 2 * m + 2 * j;
            for (size_t i = 0; i < n; ++i) mus[j * n * H + h * n + i] = b[0] + ths[(size_t)h * n + i] * b[1];  // X = [1, theta], :58-66
        }
    TRYC(gpirt_set_data_f64(T.h, ys.data()));
    TRYC(gpirt_set_priors(T.h, sds3.data(), nullptr));
    TRYC(gpirt_set_state(T.h, ths.data(), a->beta, fs.data(), a->thresholds, nullptr));
    TRYC(gpirt_set_mean_rows(T.h, mus.data()));
    TRYC(gpirt_set_iter(T.h, 0));
    TRYC(gpirt_draw_f(T.h, nullptr));
    TRYC(gpirt_get_state(T.h, nullptr, nullptr, fs.data(), nullptr, nullptr));
    const double lo = *std::min_element(ths.begin(), ths.end()), hi = *std::max_element(ths.begin(), ths.end());
    const double delta = (hi - lo) / double(P - 1);
    for (int i = 0; i + 1 < P; ++i) tind[i] = lo + double(i) * delta;
    tind[P - 1] = hi;
    TRYC(gpirt_interp1(a->device, ths.data(), (int64_t)(n * H), fs.data(), (int64_t)m, tind.data(), P, find.data()));
    TRYC(gpirt_set_data_f64(I.h, y_i.data()));
    TRYC(gpirt_set_priors(I.h, sds3.data(), nullptr));
    TRYC(gpirt_set_state(I.h, tind.data(), a->beta, find.data(), nullptr, nullptr));  // mu* of session 0 (src/draw-fstar.cpp:114)
    TRYC(gpirt_set_iter(I.h, 0));
    TRYC(gpirt_draw_fstar(I.h, nullptr));
    TRYC(gpirt_get_state(I.h, nullptr, nullptr, nullptr, nullptr, fstar.data()));
    for (int h = 0; h < H; ++h) std::memcpy(a->fstar + (size_t)h * N * m, fstar.data(), N * m * 8);  // :121-124
    return GPIRT_OK;
}
