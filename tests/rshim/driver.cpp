// TEST INFRASTRUCTURE. Drives the R-package shim (integration/r-package/src/{gpirtMCMC,recover_fstar}.cpp) without R: the shim is
// compiled against the stand-in Rcpp / Armadillo headers of oracle/refshim and linked to libgpirt_b200.so; this file plays the role of
// RcppExports.cpp (src/RcppExports.cpp:16-58 in the reference) with plain pointers instead of SEXPs. R's RNG is a caller-supplied stream.
#include <RcppArmadillo.h>

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>

Rcpp::List gpirtMCMC(const arma::cube& y, arma::mat theta, const int sample_iterations, const int burn_iterations, const int THIN,
                     const arma::mat& beta_prior_means, const arma::mat& beta_prior_sds, const arma::mat& theta_prior_means,
                     const arma::mat& theta_prior_sds, const double& theta_os, const double& theta_ls, const std::string& KERNEL,
                     arma::cube thresholds, const int constant_IRF, const bool store_f, const bool store_fstar);
Rcpp::List recover_fstar(int seed, arma::cube f, const arma::cube& y, const arma::mat& theta, const arma::cube& beta,
                         const arma::cube& thresholds, const arma::mat& beta_prior_means, const arma::mat& beta_prior_sds,
                         const int constant_IRF);

namespace {
const double* g_normals = nullptr;
uint64_t g_n_normals = 0, g_next = 0;
double g_unif = 0.5;
std::string g_err;
arma::mat mk_mat(const double* p, size_t r, size_t c) { arma::mat m(r, c); if (r * c) std::memcpy(m.memptr(), p, r * c * 8); return m; }
arma::cube mk_cube(const double* p, size_t r, size_t c, size_t s) { arma::cube q(r, c, s); if (r * c * s) std::memcpy(q.memptr(), p, r * c * s * 8); return q; }
}  // namespace

extern "C" {
double refshim_next_normal(void) {
    if (g_next >= g_n_normals) throw std::runtime_error("rshim driver: R::rnorm stream exhausted");
    return g_normals[g_next++];
}
double refshim_next_unif(void) { return g_unif; }
const char* shim_last_error(void) { return g_err.c_str(); }
uint64_t shim_normals_used(void) { return g_next; }

int shim_gpirtMCMC(uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* y, const double* theta, int sample_iterations, int burn_iterations,
                   int THIN, const double* bpm, const double* bps, const double* tpm, const double* tps, double os, double ls, const char* kernel,
                   const double* thr, int constant_IRF, const double* normals, uint64_t n_normals, double unif, double* theta_draws,
                   double* beta_draws, double* thr_draws, double* ll_draws, double* f_draws, double* fstar_draws) {
    try {
        g_normals = normals; g_n_normals = n_normals; g_next = 0; g_unif = unif;
        Rcpp::List res = gpirtMCMC(mk_cube(y, n, m, H), mk_mat(theta, n, H), sample_iterations, burn_iterations, THIN, mk_mat(bpm, 3, m),
                                   mk_mat(bps, 3, m), mk_mat(tpm, 2, n), mk_mat(tps, 2, n), os, ls, std::string(kernel), mk_cube(thr, m, C + 1, H),
                                   constant_IRF, f_draws != nullptr, fstar_draws != nullptr);
        const int n_samples = sample_iterations / THIN;
        std::memcpy(theta_draws, res["theta"].cube->memptr(), size_t(n_samples) * n * H * 8);
        std::memcpy(ll_draws, res["ll"].vec->memptr(), size_t(n_samples) * 8);
        auto unpack = [&](const char* key, double* out, size_t per) {
            if (!out) return;
            const Rcpp::RObject& l = res[key];
            for (int s = 0; s < n_samples; ++s) std::memcpy(out + size_t(s) * per, (*l.items)[s].cube->memptr(), per * 8);
        };
        unpack("beta", beta_draws, 3 * m * H);
        unpack("threshold", thr_draws, m * (C + 1) * H);
        unpack("f", f_draws, n * m * H);
        unpack("fstar", fstar_draws, 1001 * m * H);
        if (!f_draws && res["f"].kind != Rcpp::RObject::NIL) throw std::runtime_error("f should be NULL when store_f is false");
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

int shim_recover_fstar(int seed, uint64_t n, uint64_t m, uint64_t H, uint64_t C, const double* f, const double* y, const double* theta,
                       const double* beta2, const double* thr, const double* bpm, const double* bps, int constant_IRF, double* fstar) {
    try {
        Rcpp::List res = recover_fstar(seed, mk_cube(f, n, m, H), mk_cube(y, n, m, H), mk_mat(theta, n, H), mk_cube(beta2, 2, m, H),
                                       mk_cube(thr, m, C + 1, H), mk_mat(bpm, 2, m), mk_mat(bps, 2, m), constant_IRF);
        std::memcpy(fstar, res["fstar"].cube->memptr(), size_t(1001) * m * H * 8);
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return 1; }
}
}  // extern "C"
