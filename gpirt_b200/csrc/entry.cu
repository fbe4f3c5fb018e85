// The two Rcpp entry points of the reference as C-ABI calls with host buffers:
//   gpirt_mcmc            <- gpirtMCMC()      src/gpirtMCMC.cpp:25-406
//   gpirt_recover_fstar   <- recover_fstar()  src/recover_fstar.cpp:8-93
// Host side only: sample storage, thinning and the progress/interrupt poll; all arithmetic runs in the
// kernels behind the subsystem calls.
#include <cstdio>
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/gpirt_b200.h"

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
        if (data_from) TRYC(gpirt_copy_data(h, data_from));
        else TRYC(gpirt_set_data_f64(h, a->y));
        TRYC(gpirt_set_priors(h, a->beta_prior_sds, a->theta_prior_sds));
        TRYC(gpirt_set_state(h, a->theta, a->beta_init, a->f_init, a->thresholds, nullptr));
        TRYC(gpirt_set_iter(h, 0xFFFFFFFFu));  // initial draws use a sweep id no sweep reaches
        TRYC(gpirt_init_draw(h, a->beta_prior_means, a->beta_init == nullptr, a->f_init == nullptr && a->f_init_z == nullptr));
        if (!a->f_init && a->f_init_z) TRYC(gpirt_init_f_from_z(h, a->f_init_z));
        TRYC(gpirt_set_iter(h, 0));
        n_samples = a->sample_iterations / a->THIN;  // src/gpirtMCMC.cpp:48
        (void)n_chains;
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

extern "C" int gpirt_mcmc(const gpirt_mcmc_args* a) {
    if (!args_ok(a)) return GPIRT_ERR_INVALID;
    return gpirt_mcmc_chains(a, 1);
}

extern "C" int gpirt_recover_fstar(const gpirt_recover_fstar_args* a) {
    if (!a || !a->f || !a->y || !a->theta || !a->beta || !a->thresholds || !a->beta_prior_sds || !a->fstar) return GPIRT_ERR_INVALID;
    gpirt_config cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.n = a->n; cfg.m_total = a->m; cfg.horizon = a->horizon; cfg.C = a->C; cfg.constant_IRF = a->constant_IRF; cfg.p_beta = 2;
    cfg.theta_os = 1.0; cfg.theta_ls = 10.0; cfg.kernel = 0; cfg.device = a->device;
    cfg.seed = (uint64_t)(uint32_t)a->seed;  // ws_pool.seed_all(seed), src/recover_fstar.cpp:80
    cfg.chain = 0; cfg.rank = 0; cfg.world = 1;
    Guard g;
    TRYC(gpirt_create(&cfg, &g.h));
    gpirt_sampler* h = g.h;
    TRYC(gpirt_set_data_f64(h, a->y));
    // recover_fstar passes a 2 x m beta_prior_sds (R/recover_fstar.R:37-38); K() reads rows 0..2 of column 0
    // (src/covariance-function.cpp:12-14) — an out-of-bounds read in the reference. Row 2 is taken as 0 here.
    std::vector<double> sds3(3 * (size_t)a->m, 0.0);
    for (int j = 0; j < a->m; ++j) { sds3[3 * (size_t)j] = a->beta_prior_sds[2 * (size_t)j]; sds3[3 * (size_t)j + 1] = a->beta_prior_sds[2 * (size_t)j + 1]; }
    TRYC(gpirt_set_priors(h, sds3.data(), nullptr));
    TRYC(gpirt_set_state(h, a->theta, a->beta, a->f, a->thresholds, nullptr));
    TRYC(gpirt_set_iter(h, 0));
    TRYC(gpirt_draw_f(h, nullptr));      // src/recover_fstar.cpp:87-88
    TRYC(gpirt_draw_fstar(h, nullptr));  // :89-90
    return gpirt_get_state(h, nullptr, nullptr, nullptr, nullptr, a->fstar);
}
