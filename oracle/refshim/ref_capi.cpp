// TEST INFRASTRUCTURE — C entry points into the reference's own translation units (compiled from
// /root/reference/src against the stand-in headers of this directory; see oracle/Makefile `ref`).
// tests/test_oracle_vs_ref.py drives these to pin the CPU oracle (oracle/orc_*.{h,cpp}) to the reference's code;
// tests/golden/make_golden.py records their outputs as fixtures that travel to the GPU box.
// All matrices column-major float64 as Armadillo lays them out; y uses NaN for missing.
#include <omp.h>

#include <cstdint>
#include <cstring>
#include <string>

#include "gpirt.h"  // the reference's header (-I/root/reference/src)

// the two Rcpp entry points (src/gpirtMCMC.cpp:25, src/recover_fstar.cpp:8) have no prototype in gpirt.h
Rcpp::List gpirtMCMC(const arma::cube& y, arma::mat theta, const int sample_iterations, const int burn_iterations, const int THIN,
                     const arma::mat& beta_prior_means, const arma::mat& beta_prior_sds, const arma::mat& theta_prior_means,
                     const arma::mat& theta_prior_sds, const double& theta_os, const double& theta_ls, const std::string& KERNEL,
                     arma::cube thresholds, const int constant_IRF, const bool store_f, const bool store_fstar);
Rcpp::List recover_fstar(int seed, arma::cube f, const arma::cube& y, const arma::mat& theta, const arma::cube& beta,
                         const arma::cube& thresholds, const arma::mat& beta_prior_means, const arma::mat& beta_prior_sds,
                         const int constant_IRF);

namespace {
const double* g_normals = nullptr;
size_t g_n_normals = 0, g_next = 0;
thread_local std::string g_err;

arma::vec mk_vec(const double* p, size_t n) { arma::vec v(n); std::memcpy(v.memptr(), p, n * 8); return v; }
arma::mat mk_mat(const double* p, size_t r, size_t c) { arma::mat m(r, c); if (r * c) std::memcpy(m.memptr(), p, r * c * 8); return m; }
arma::cube mk_cube(const double* p, size_t r, size_t c, size_t s) { arma::cube q(r, c, s); if (r * c * s) std::memcpy(q.memptr(), p, r * c * s * 8); return q; }
void put(double* out, const arma::cube& c) { if (out) std::memcpy(out, c.memptr(), size_t(c.n_elem) * 8); }

struct Masks {
    arma::field<arma::uvec> obs_items, obs_persons, obs_persons_combined;
    Masks(const arma::cube& y) : obs_items(y.n_rows, y.n_slices), obs_persons(y.n_cols, y.n_slices), obs_persons_combined(y.n_cols, 1) {
        const arma::uword n = y.n_rows, m = y.n_cols, H = y.n_slices;  // as src/gpirtMCMC.cpp:93-125
        for (arma::uword h = 0; h < H; ++h) {
            for (arma::uword i = 0; i < n; ++i) obs_items(i, h) = arma::find_finite(y.slice(h).row(i));
            for (arma::uword j = 0; j < m; ++j) obs_persons(j, h) = arma::find_finite(y.slice(h).col(j));
        }
        for (arma::uword j = 0; j < m; ++j) {
            arma::uword total = 0;
            for (arma::uword h = 0; h < H; ++h) total += obs_persons(j, h).n_elem;
            arma::uvec all(total);
            arma::uword pos = 0;
            for (arma::uword h = 0; h < H; ++h)
                for (arma::uword k = 0; k < obs_persons(j, h).n_elem; ++k) all(pos++) = obs_persons(j, h)(k) + h * n;
            obs_persons_combined(j, 0) = all;
        }
    }
};
}  // namespace

#define GUARD(body)                                   \
    try { body; return 0; }                           \
    catch (const std::exception& e) { g_err = e.what(); return 1; } \
    catch (...) { g_err = "unknown exception"; return 2; }

extern "C" {

double refshim_next_normal(void) {
    if (g_next >= g_n_normals) throw std::runtime_error("refshim: R::rnorm stream exhausted");
    return g_normals[g_next++];
}
void refshim_set_normals(const double* z, uint64_t n) { g_normals = z; g_n_normals = n; g_next = 0; }
uint64_t refshim_normals_used(void) { return g_next; }
const char* ref_last_error(void) { return g_err.c_str(); }
void ref_set_threads(int k) { omp_set_num_threads(k); }
int ref_get_threads(void) { return get_num_threads(); }

double ref_pnorm(double x) { return R::pnorm(x, 0, 1, 1, 0); }

int ref_K(const double* x1, uint64_t n1, const double* x2, uint64_t n2, const double* sds3, double* out) {
    GUARD(arma::mat r = K(mk_vec(x1, n1), mk_vec(x2, n2), mk_vec(sds3, 3)); std::memcpy(out, r.memptr(), n1 * n2 * 8));
}
int ref_K_time(const double* x1, uint64_t n1, const double* x2, uint64_t n2, double os, double ls, const double* tsds2, const char* kernel,
               double* out) {
    GUARD(arma::mat r = K_time(mk_vec(x1, n1), mk_vec(x2, n2), os, ls, mk_vec(tsds2, 2), std::string(kernel));
          std::memcpy(out, r.memptr(), n1 * n2 * 8));
}
int ref_ll_bar_sparse(const double* f, const double* y, const double* mu, const double* thr, uint64_t n, uint64_t C, double* out) {
    GUARD(arma::vec yv = mk_vec(y, n); arma::uvec obs = arma::find_finite(yv.head(n));
          *out = ll_bar_sparse(mk_vec(f, n), yv, mk_vec(mu, n), mk_vec(thr, C + 1), obs));
}
int ref_delta_to_threshold(const double* d, uint64_t C, double* thr) {
    GUARD(arma::vec t = delta_to_threshold(mk_vec(d, C - 1)); std::memcpy(thr, t.memptr(), (C + 1) * 8));
}
int ref_threshold_to_delta(const double* thr, uint64_t C, double* d) {
    GUARD(arma::vec t = threshold_to_delta(mk_vec(thr, C + 1)); std::memcpy(d, t.memptr(), (C - 1) * 8));
}
// update_cholesky_cache (src/cholesky-cache.cpp:11-40): L is n x n x H, L_time H x H x n (may be NULL)
int ref_update_cholesky_cache(const double* theta, uint64_t n, uint64_t H, const double* sds, uint64_t m, double os, double ls,
                              const char* kernel, double* L, double* L_time) {
    GUARD(CholeskyCache cache(n, H); update_cholesky_cache(cache, mk_mat(theta, n, H), mk_mat(sds, 3, m), os, ls, std::string(kernel));
          put(L, cache.L); put(L_time, cache.L_time));
}
int ref_double_solve(const double* L, uint64_t n, const double* X, uint64_t nrhs, double* out) {
    GUARD(arma::mat r = double_solve(mk_mat(L, n, n), mk_mat(X, n, nrhs)); std::memcpy(out, r.memptr(), n * nrhs * 8));
}

// ---- per-block samplers; each call builds a fresh WorkspacePool and seeds it with seed_all(seed) ----
int ref_draw_f(uint32_t seed, uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* f, const double* theta, const double* y,
               const double* sds, const double* mu, const double* thr, int constant_IRF, double* f_new) {
    GUARD(arma::cube yc = mk_cube(y, n, m, H); Masks mk(yc); CholeskyCache cache(n, H); WorkspacePool pool(n, m, H, get_num_threads());
          pool.seed_all(seed); arma::cube out(n, m, H);
          draw_f(out, mk_cube(f, n, m, H), mk_mat(theta, n, H), yc, cache, mk_mat(sds, 3, m), mk_cube(mu, n, m, H), mk_cube(thr, m, C + 1, H),
                 constant_IRF, mk.obs_persons, mk.obs_persons_combined, pool);
          put(f_new, out));
}
int ref_draw_fstar(uint32_t seed, uint64_t n, uint64_t m, uint64_t H, const double* f, const double* theta, const double* sds,
                   const double* mu_star, int constant_IRF, double* fstar) {
    GUARD(CholeskyCache cache(n, H); arma::mat th = mk_mat(theta, n, H); arma::mat s = mk_mat(sds, 3, m);
          update_cholesky_cache(cache, th, s, 0, 0, ""); WorkspacePool pool(n, m, H, get_num_threads()); pool.seed_all(seed);
          arma::vec tstar = arma::regspace<arma::vec>(-5.0, 0.01, 5.0); arma::cube out(tstar.n_elem, m, H);
          draw_fstar(out, mk_cube(f, n, m, H), th, tstar, s, cache, mk_cube(mu_star, tstar.n_elem, m, H), constant_IRF, pool);
          put(fstar, out));
}
int ref_draw_theta(uint32_t seed, uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* y, const double* theta, const double* tps,
                   const double* fstar, const double* mu_star, const double* thr, double os, double ls, const char* kernel, const double* sds,
                   double* theta_new) {
    GUARD(arma::cube yc = mk_cube(y, n, m, H); Masks mk(yc); CholeskyCache cache(n, H); arma::mat th = mk_mat(theta, n, H);
          update_cholesky_cache(cache, th, mk_mat(sds, 3, m), os, ls, std::string(kernel));
          WorkspacePool pool(n, m, H, get_num_threads()); pool.seed_all(seed);
          arma::vec tstar = arma::regspace<arma::vec>(-5.0, 0.01, 5.0); arma::mat out(n, H);
          draw_theta(out, tstar, yc, th, mk_mat(tps, 2, n), mk_cube(fstar, 1001, m, H), mk_cube(mu_star, 1001, m, H), mk_cube(thr, m, C + 1, H), os,
                     ls, std::string(kernel), mk.obs_items, cache, pool);
          std::memcpy(theta_new, out.memptr(), n * H * 8));
}
int ref_draw_beta(uint32_t seed, uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* beta, const double* theta, const double* y,
                  const double* f, const double* prior_means, const double* prior_sds, const double* thr, double* beta_new) {
    GUARD(arma::cube yc = mk_cube(y, n, m, H); Masks mk(yc); arma::mat th = mk_mat(theta, n, H);
          arma::cube X(n, 3, H); X.col(0) = arma::ones<arma::mat>(n, H); X.col(1) = th; X.col(2) = arma::pow(th, 2);  // src/gpirtMCMC.cpp:138-141
          WorkspacePool pool(n, m, H, get_num_threads()); pool.seed_all(seed); arma::cube out(3, m, H);
          draw_beta(out, mk_cube(beta, 3, m, H), X, yc, mk_cube(f, n, m, H), mk_mat(prior_means, 3, m), mk_mat(prior_sds, 3, m),
                    mk_cube(thr, m, C + 1, H), mk.obs_persons, pool);
          put(beta_new, out));
}
int ref_draw_threshold(uint32_t seed, uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* thr, const double* y, const double* f,
                       const double* mu, int constant_IRF, double* thr_new) {
    GUARD(arma::cube yc = mk_cube(y, n, m, H); Masks mk(yc); WorkspacePool pool(n, m, H, get_num_threads()); pool.seed_all(seed);
          arma::cube out(m, C + 1, H);
          draw_threshold(out, mk_cube(thr, m, C + 1, H), yc, mk_cube(f, n, m, H), mk_cube(mu, n, m, H), constant_IRF, mk.obs_persons,
                         mk.obs_persons_combined, pool);
          put(thr_new, out));
}

// ---- the two entry points ----
// Outputs: theta_draws n_samples x n x H; beta_draws 3 x m x H x n_samples; thr_draws m x (C+1) x H x n_samples; ll n_samples;
// f_draws n x m x H x n_samples and fstar_draws 1001 x m x H x n_samples when non-NULL. init_normals feeds R::rnorm
// (3m draws for beta0, then n per item for f0 = L z: src/gpirtMCMC.cpp:149-163).
int ref_gpirtMCMC(uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* y, const double* theta, int sample_iterations, int burn_iterations,
                  int THIN, const double* bpm, const double* bps, const double* tpm, const double* tps, double os, double ls, const char* kernel,
                  const double* thr, int constant_IRF, const double* init_normals, uint64_t n_init, double* theta_draws, double* beta_draws,
                  double* thr_draws, double* ll_draws, double* f_draws, double* fstar_draws) {
    GUARD(
        refshim_set_normals(init_normals, n_init);
        Rcpp::List res = gpirtMCMC(mk_cube(y, n, m, H), mk_mat(theta, n, H), sample_iterations, burn_iterations, THIN, mk_mat(bpm, 3, m),
                                   mk_mat(bps, 3, m), mk_mat(tpm, 2, n), mk_mat(tps, 2, n), os, ls, std::string(kernel), mk_cube(thr, m, C + 1, H),
                                   constant_IRF, f_draws != nullptr, fstar_draws != nullptr);
        const int n_samples = sample_iterations / THIN;
        put(theta_draws, *res["theta"].cube);
        std::memcpy(ll_draws, res["ll"].vec->memptr(), size_t(n_samples) * 8);
        auto unpack = [&](const char* key, double* out, size_t per) {
            if (!out) return;
            const Rcpp::RObject& l = res[key];
            for (int s = 0; s < n_samples; ++s) std::memcpy(out + size_t(s) * per, (*l.items)[s].cube->memptr(), per * 8);
        };
        unpack("beta", beta_draws, 3 * m * H);
        unpack("threshold", thr_draws, m * (C + 1) * H);
        unpack("f", f_draws, n * m * H);
        unpack("fstar", fstar_draws, 1001 * m * H));
}
int ref_recover_fstar(int seed, uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* f, const double* y, const double* theta,
                      const double* beta2, const double* thr, const double* bpm, const double* bps, uint64_t bps_rows, int constant_IRF,
                      double* fstar) {
    GUARD(Rcpp::List res = recover_fstar(seed, mk_cube(f, n, m, H), mk_cube(y, n, m, H), mk_mat(theta, n, H), mk_cube(beta2, 2, m, H),
                                         mk_cube(thr, m, C + 1, H), mk_mat(bpm, bps_rows, m), mk_mat(bps, bps_rows, m), constant_IRF);
          put(fstar, *res["fstar"].cube));
}

}  // extern "C"
