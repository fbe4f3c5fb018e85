// TEST INFRASTRUCTURE — flat C entry points over the CPU oracle, loaded with ctypes by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs only.
#include <omp.h>

#include <cstdio>
#include <cstring>

#include "orc_sampler.h"

using namespace orc;

extern "C" {

struct orc_src {
    int32_t mode;  // 0 supplied, 1 philox, 2 reference mt19937 ThreadRNG
    const double* z; uint64_t zstride;
    const double* u; uint64_t ustride;
    uint64_t seed; uint32_t chain, iter;
    uint32_t session, pad_;
};

static Src to_src(Sampler* S, const orc_src* c, int block) {
    static const uint32_t zs[N_BLOCKS] = {ST_F_Z, ST_FSTAR_Z, ST_THETA_Z, ST_BETA_Z, ST_THR_Z};
    static const uint32_t us[N_BLOCKS] = {ST_F_U, ST_F_U /*unused*/, ST_THETA_U, ST_BETA_U, ST_THR_U};
    Src s;
    s.mode = c->mode; s.z = c->z; s.zstride = c->zstride; s.u = c->u; s.ustride = c->ustride;
    s.seed = c->seed; s.chain = c->chain; s.iter = c->iter; s.zs = zs[block]; s.us = us[block];
    s.session = c->session;
    s.mts = S->mts_shared ? S->mts_shared : &S->mts;
    if (S->record) {
        const size_t units = block == B_THETA ? S->n : S->m;
        S->rec_z[block].assign(units, {});
        S->rec_u[block].assign(units, {});
        s.rec_z = &S->rec_z[block];
        s.rec_u = &S->rec_u[block];
    }
    return s;
}

// ---- optional BLAS backend of the TIMED baseline (orc_blas.h); the parity tests leave it off ----
int orc_blas_load(const char* path, int threads_all) { return blas_load(path, threads_all); }
void orc_blas_off(void) { blas_off(); }
int orc_blas_on(void) { return blas().on ? 1 : 0; }

// ---- leaves -------------------------------------------------------------------------------
double orc_pnorm(double x) { return pnorm(x); }
void orc_pnorm_vec(const double* x, uint64_t n, double* out) { for (uint64_t i = 0; i < n; ++i) out[i] = pnorm(x[i]); }
void orc_K(const double* x1, uint64_t n1, const double* x2, uint64_t n2, const double* sds, double* out) { K(x1, n1, x2, n2, sds, out); }
void orc_K_time(const double* x1, uint64_t n1, const double* x2, uint64_t n2, double os, double ls, const double* tsds, int kernel, double* out) {
    K_time(x1, n1, x2, n2, os, ls, tsds, kernel, out);
}
int orc_chol(double* A, uint64_t n) { return chol_lower(A, n); }
int orc_chol_blocked(double* A, uint64_t n) { return chol_lower_blocked(A, n); }
int orc_chol_ld(double* A, uint64_t n) {  // long-double factorisation, rounded back to double
    std::vector<long double> B(A, A + n * n);
    int info = chol_lower(B.data(), n);
    for (uint64_t i = 0; i < n * n; ++i) A[i] = double(B[i]);
    return info;
}
void orc_double_solve(const double* L, uint64_t n, double* x) { double_solve(L, n, x); }
void orc_trsv_lower(const double* L, uint64_t n, double* x) { trsv_lower(L, n, x); }
double orc_ll_bar_sparse(const double* f, const double* y, const double* mu, const double* thr, uint64_t n) {
    std::vector<uint64_t> obs;
    for (uint64_t i = 0; i < n; ++i) if (std::isfinite(y[i])) obs.push_back(i);  // find_finite, src/gpirtMCMC.cpp:108
    return ll_bar_sparse(f, y, mu, thr, obs.data(), obs.size());
}
void orc_delta_to_threshold(const double* d, uint64_t C, double* thr) { delta_to_threshold(d, C, thr); }
void orc_threshold_to_delta(const double* thr, uint64_t C, double* d) { threshold_to_delta(thr, C, d); }
int orc_grid_index(double theta) { return grid_index(theta); }
double orc_grid_value(int idx) { return grid_value(idx); }

void orc_philox_uniform(uint64_t seed, uint32_t chain, uint32_t iter, uint32_t stream, uint32_t unit, uint64_t count, double* out) {
    for (uint64_t k = 0; k < count; k += 2) {
        double a, b;
        Philox::uniform2(seed, chain, iter, stream, unit, uint32_t(k >> 1), a, b);
        out[k] = a;
        if (k + 1 < count) out[k + 1] = b;
    }
}
void orc_philox_normal(uint64_t seed, uint32_t chain, uint32_t iter, uint32_t stream, uint32_t unit, uint64_t count, double* out) {
    for (uint64_t k = 0; k < count; k += 2) {
        double a, b;
        Philox::normal2(seed, chain, iter, stream, unit, uint32_t(k >> 1), a, b);
        out[k] = a;
        if (k + 1 < count) out[k + 1] = b;
    }
}
void orc_philox_raw(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out) {
    uint32_t c[4] = {c0, c1, c2, c3};
    Philox::run(c, k0, k1);
    std::memcpy(out, c, sizeof(c));
}

// ---- sampler ------------------------------------------------------------------------------
void* orc_create(uint64_t n, uint64_t m, uint64_t C, const double* y, const double* sds, const double* tps) {
    Sampler* S = new Sampler();
    S->init(n, m, C, y, sds, tps);
    return S;
}
void orc_destroy(void* h) { delete static_cast<Sampler*>(h); }
void orc_set_opts(void* h, int faithful, int p_beta, int theta_regime, double theta_os) {
    Sampler* S = static_cast<Sampler*>(h);
    S->faithful = faithful; S->p_beta = p_beta; S->theta_regime = theta_regime; S->theta_os = theta_os;
}
void orc_set_state(void* h, const double* theta, const double* beta, const double* f, const double* thr, const double* fstar) {
    Sampler* S = static_cast<Sampler*>(h);
    if (theta) S->theta.assign(theta, theta + S->n);
    if (beta) S->beta.assign(beta, beta + 3 * S->m);
    if (f) S->f.assign(f, f + S->n * S->m);
    if (thr) S->thr.assign(thr, thr + S->m * (S->C + 1));
    if (fstar) S->fstar.assign(fstar, fstar + NGRID * S->m);
}
void orc_get_state(void* h, double* theta, double* beta, double* f, double* thr, double* fstar) {
    Sampler* S = static_cast<Sampler*>(h);
    if (theta) std::memcpy(theta, S->theta.data(), S->n * 8);
    if (beta) std::memcpy(beta, S->beta.data(), 3 * S->m * 8);
    if (f) std::memcpy(f, S->f.data(), S->n * S->m * 8);
    if (thr) std::memcpy(thr, S->thr.data(), S->m * (S->C + 1) * 8);
    if (fstar) std::memcpy(fstar, S->fstar.data(), NGRID * S->m * 8);
}
void orc_get_L(void* h, double* L) { Sampler* S = static_cast<Sampler*>(h); std::memcpy(L, S->L.data(), S->n * S->n * 8); }
void orc_set_L(void* h, const double* L) { Sampler* S = static_cast<Sampler*>(h); S->L.assign(L, L + S->n * S->n); }
void orc_get_mu(void* h, double* mu, double* mu_star) {
    Sampler* S = static_cast<Sampler*>(h);
    if (mu) std::memcpy(mu, S->mu.data(), S->n * S->m * 8);
    if (mu_star) std::memcpy(mu_star, S->mu_star.data(), NGRID * S->m * 8);
}
void orc_compute_mu(void* h) { static_cast<Sampler*>(h)->compute_mu(); }
int orc_update_cholesky(void* h) { return static_cast<Sampler*>(h)->update_cholesky(); }
void orc_mt_seed_all(void* h, uint32_t base) {  // WorkspacePool::seed_all, src/gpirt.h:178-182
    Sampler* S = static_cast<Sampler*>(h);
    for (size_t t = 0; t < S->mts.size(); ++t) S->mts[t].seed(base + uint32_t(t) * 1000u);
}
// worker threads of samplers created AFTER this call (single-thread runs reproduce the reference's RNG order)
void orc_set_threads(int k) {
    omp_set_num_threads(k);
    if (blas().on) { blas().threads_all = k; blas().set_threads(k); }
}
void orc_record(void* h, int on) { static_cast<Sampler*>(h)->record = on != 0; }
// draws of the last call of `block`: z_out is zcap x units, u_out is ucap x units (column-major, zero padded); counts per unit
int orc_get_record(void* h, int block, double* z_out, uint64_t zcap, double* u_out, uint64_t ucap, int32_t* nz, int32_t* nu) {
    Sampler* S = static_cast<Sampler*>(h);
    const auto& Z = S->rec_z[block];
    const auto& U = S->rec_u[block];
    int over = 0;
    for (size_t k = 0; k < Z.size(); ++k) {
        nz[k] = int32_t(Z[k].size()); nu[k] = int32_t(U[k].size());
        if (Z[k].size() > zcap || U[k].size() > ucap) { over = 1; continue; }
        for (size_t i = 0; i < Z[k].size(); ++i) z_out[k * zcap + i] = Z[k][i];
        for (size_t i = 0; i < U[k].size(); ++i) u_out[k * ucap + i] = U[k][i];
    }
    return over;
}
int orc_num_threads(void* h) { return int(static_cast<Sampler*>(h)->mts.size()); }

void orc_draw_f(void* h, const orc_src* c, const double* nu) { Sampler* S = static_cast<Sampler*>(h); S->draw_f(to_src(S, c, B_F), nu); }
int orc_draw_fstar(void* h, const orc_src* c) { Sampler* S = static_cast<Sampler*>(h); return S->draw_fstar(to_src(S, c, B_FSTAR)); }
// horizon > 1: session `dst` consumes the ThreadRNGs of `src` (one ws_pool for all sessions, src/gpirtMCMC.cpp:90)
int orc_draw_fstar_cst(void* h, const orc_src* c) { Sampler* S = static_cast<Sampler*>(h); return S->draw_fstar_cst(to_src(S, c, B_FSTAR)); }
// constant-IRF panel: the stacked sampler's mean is assembled from the sessions' own mu = X_h beta_h (src/draw-f.cpp:93-104)
void orc_set_mu(void* h, const double* mu) { Sampler* S = static_cast<Sampler*>(h); S->mu.assign(mu, mu + S->n * S->m); }
void orc_share_mts(void* dst, void* src) { static_cast<Sampler*>(dst)->mts_shared = &static_cast<Sampler*>(src)->mts; }
void orc_panel_draw_theta(void** hs, uint64_t H, int regime, const double* L_time, const orc_src* cs) {
    std::vector<Sampler*> S(H);
    std::vector<Src> srcs;
    for (uint64_t h = 0; h < H; ++h) S[h] = static_cast<Sampler*>(hs[h]);
    if (regime == 1) for (uint64_t h = 0; h < H; ++h) srcs.push_back(to_src(S[h], &cs[h], B_THETA));
    else srcs.push_back(to_src(S[0], &cs[0], B_THETA));
    panel_draw_theta(S.data(), H, regime, L_time, srcs.data());
}
void orc_draw_theta(void* h, const orc_src* c) { Sampler* S = static_cast<Sampler*>(h); S->draw_theta(to_src(S, c, B_THETA)); }
void orc_gather_f(void* h) { static_cast<Sampler*>(h)->gather_f(); }
void orc_draw_beta(void* h, const orc_src* c) { Sampler* S = static_cast<Sampler*>(h); S->draw_beta(to_src(S, c, B_BETA)); }
void orc_draw_threshold(void* h, const orc_src* c) { Sampler* S = static_cast<Sampler*>(h); S->draw_threshold(to_src(S, c, B_THR)); }
double orc_total_ll(void* h) { return static_cast<Sampler*>(h)->total_ll(); }
void orc_item_ll(void* h, double* out) {  // per-item ll_bar_sparse with the current state
    Sampler* S = static_cast<Sampler*>(h);
    std::vector<double> t(S->C + 1);
    for (size_t j = 0; j < S->m; ++j) {
        for (size_t c = 0; c <= S->C; ++c) t[c] = S->thr[j + c * S->m];
        out[j] = ll_bar_sparse(&S->f[j * S->n], &S->y[j * S->n], &S->mu[j * S->n], t.data(), S->obs_persons[j].data(), S->obs_persons[j].size());
    }
}
int orc_sweep(void* h, const orc_src* c /* N_BLOCKS entries */) {
    Sampler* S = static_cast<Sampler*>(h);
    Src s[N_BLOCKS];
    for (int b = 0; b < N_BLOCKS; ++b) s[b] = to_src(S, &c[b], b);
    return S->sweep(s);
}
void orc_get_times(void* h, double* fixed, double* items, int reset) {
    Sampler* S = static_cast<Sampler*>(h);
    *fixed = S->t_fixed; *items = S->t_items;
    if (reset) S->t_fixed = S->t_items = 0.0;
}
double orc_last_ll(void* h) { return static_cast<Sampler*>(h)->ll; }
int orc_get_diag(void* h, int block, int32_t* evals, double* margin) {
    Sampler* S = static_cast<Sampler*>(h);
    const Diag& d = S->diag[block];
    if (evals) std::memcpy(evals, d.evals.data(), d.evals.size() * 4);
    if (margin) std::memcpy(margin, d.margin.data(), d.margin.size() * 8);
    return d.overrun ? 1 : 0;
}

// The reference's sweep loop with its own RNG discipline (src/gpirtMCMC.cpp:261-356): sweeps
// iter0 .. iter0+iters-1, each reseeding every worker's mt19937 with iter*10000 + tid*1000.
// Stored draws: theta (n per sweep), ll (1 per sweep); caller may pass NULL.
int orc_run_mt(void* h, int iter0, int iters, double* theta_out, double* ll_out) {
    Sampler* S = static_cast<Sampler*>(h);
    orc_src c[N_BLOCKS];
    std::memset(c, 0, sizeof(c));
    for (int b = 0; b < N_BLOCKS; ++b) c[b].mode = 2;
    for (int it = 0; it < iters; ++it) {
        orc_mt_seed_all(h, uint32_t((iter0 + it) * 10000));
        int info = orc_sweep(h, c);
        if (info) return info;
        if (theta_out) std::memcpy(theta_out + size_t(it) * S->n, S->theta.data(), S->n * 8);
        if (ll_out) ll_out[it] = S->ll;
    }
    return 0;
}

}  // extern "C"
