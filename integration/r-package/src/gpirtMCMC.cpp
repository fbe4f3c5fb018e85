// Drop-in replacement for the body of gpirtMCMC() (reference: src/gpirtMCMC.cpp:25-406) over libgpirt_b200.so.
// Same exported signature (src/gpirt.h / src/RcppExports.cpp:16-37), same returned list (src/gpirtMCMC.cpp:396-403); the generated
// RcppExports.cpp, NAMESPACE and the R code stay as they are. Initial values are drawn with R's RNG exactly as the reference does
// (src/gpirtMCMC.cpp:149-163: beta0 ~ R::rnorm per (session, item, coefficient), then z ~ N(0,1) per (session, item, respondent) for
// f0 = L z, src/mvnormal.h:7-14); the product L z runs on the device (gpirt_mcmc_args.f_init_z).
// Compile-checked and run in this repository against stand-in Rcpp/Armadillo headers (tests/rshim); builds unchanged against RcppArmadillo.
#include <RcppArmadillo.h>

#include <algorithm>
#include <string>
#include <vector>

#include "gpirt_b200.h"

namespace {
int gpirt_b200_poll(int iter, int total, void*) {  // src/gpirtMCMC.cpp:262-264
    Rprintf("\r%6.3f %% complete", 100.0 * iter / total);
    try {
        Rcpp::checkUserInterrupt();
    } catch (...) {
        return 1;  // nothing may unwind through the C library: it stops the run and returns an error code
    }
    return 0;
}
}  // namespace

// [[Rcpp::export(.gpirtMCMC)]]
Rcpp::List gpirtMCMC(const arma::cube& y, arma::mat theta, const int sample_iterations, const int burn_iterations, const int THIN,
                     const arma::mat& beta_prior_means, const arma::mat& beta_prior_sds, const arma::mat& theta_prior_means,
                     const arma::mat& theta_prior_sds, const double& theta_os, const double& theta_ls, const std::string& KERNEL,
                     arma::cube thresholds, const int constant_IRF, const bool store_f, const bool store_fstar) {
    const arma::uword n = y.n_rows, m = y.n_cols, horizon = y.n_slices, C1 = thresholds.n_cols;
    const int n_samples = int(sample_iterations / THIN);  // src/gpirtMCMC.cpp:48
    {   // the memory estimate and the > 10 000 MB warning, printed before anything is allocated (src/gpirtMCMC.cpp:47-82)
        gpirt_mcmc_args e{};
        e.n = int(n); e.m = int(m); e.horizon = int(horizon); e.C = int(C1) - 1;
        e.sample_iterations = sample_iterations; e.burn_iterations = burn_iterations; e.THIN = THIN; e.store_f = store_f; e.store_fstar = store_fstar;
        gpirt_mem_estimate est;
        char text[2048];
        if (gpirt_memory_estimate(&e, &est, text, sizeof(text)) == GPIRT_OK) Rcpp::Rcout << text;
    }

    // initial values with R's RNG, in the reference's order
    // constant_IRF == 1 draws beta for session 0 only and ONE f0 = L_0 z per item (src/gpirtMCMC.cpp:176-188): the library reads slice 0 then
    arma::cube beta0(3, m, horizon, arma::fill::zeros), z0(n, m, horizon, arma::fill::zeros);
    const arma::uword h_drawn = constant_IRF ? 1 : horizon;
    for (arma::uword h = 0; h < h_drawn; ++h)
        for (arma::uword j = 0; j < m; ++j)
            for (arma::uword p = 0; p < 3; ++p) beta0(p, j, h) = R::rnorm(beta_prior_means(p, j), beta_prior_sds(p, j));
    for (arma::uword h = 0; h < h_drawn; ++h)
        for (arma::uword j = 0; j < m; ++j)
            for (arma::uword i = 0; i < n; ++i) z0(i, j, h) = R::rnorm(0.0, 1.0);

    arma::cube theta_draws(n_samples, n, horizon);
    arma::vec ll_draws(n_samples);
    const size_t beta_sz = size_t(3) * m * horizon, thr_sz = size_t(m) * C1 * horizon, f_sz = size_t(n) * m * horizon,
                 fstar_sz = size_t(GPIRT_NGRID) * m * horizon;
    std::vector<double> beta_buf(beta_sz * n_samples), thr_buf(thr_sz * n_samples), f_buf(store_f ? f_sz * n_samples : 0),
        fstar_buf(store_fstar ? fstar_sz * n_samples : 0);

    gpirt_mcmc_args a{};
    a.y = y.memptr(); a.n = int(n); a.m = int(m); a.horizon = int(horizon);
    a.theta = theta.memptr();
    a.sample_iterations = sample_iterations; a.burn_iterations = burn_iterations; a.THIN = THIN;
    a.beta_prior_means = beta_prior_means.memptr(); a.beta_prior_sds = beta_prior_sds.memptr();
    a.theta_prior_means = theta_prior_means.memptr(); a.theta_prior_sds = theta_prior_sds.memptr();
    a.theta_os = theta_os; a.theta_ls = theta_ls; a.KERNEL = KERNEL.c_str();
    a.thresholds = thresholds.memptr(); a.C = int(C1) - 1;
    a.constant_IRF = constant_IRF; a.store_f = store_f; a.store_fstar = store_fstar;
    a.beta_init = beta0.memptr(); a.f_init = nullptr; a.f_init_z = z0.memptr();
    // The reference's sweeps use mt19937 streams seeded by the sweep number alone (src/gpirtMCMC.cpp:267); this library's
    // counter-based streams want a chain seed: one more draw from R's generator, so set.seed() governs the whole chain.
    a.seed = (uint64_t)(R::unif_rand() * 4294967296.0); a.chain = 0; a.device = 0;
    a.theta_draws = theta_draws.memptr(); a.beta_draws = beta_buf.data(); a.threshold_draws = thr_buf.data();
    a.f_draws = store_f ? f_buf.data() : nullptr; a.fstar_draws = store_fstar ? fstar_buf.data() : nullptr;
    a.ll_draws = ll_draws.memptr();
    a.progress = gpirt_b200_poll; a.user = nullptr;

    Rcpp::Rcout << "Setting up gpirtMCMC...\n";
    const int rc = gpirt_mcmc(&a);
    if (rc != GPIRT_OK) {
        Rcpp::checkUserInterrupt();        // a pending interrupt surfaces as the usual R condition
        Rcpp::stop(gpirt_last_error());    // e.g. "chol(): decomposition failed", as arma::chol would raise through END_RCPP
    }
    Rprintf("\r100.000 %% complete\n");

    // lists of cubes, as src/gpirtMCMC.cpp:359-403
    auto slices = [&](const std::vector<double>& buf, arma::uword r, arma::uword c, size_t sz) {
        Rcpp::List out(n_samples);
        for (int s = 0; s < n_samples; ++s) {
            arma::cube d(r, c, horizon);
            std::copy(buf.begin() + size_t(s) * sz, buf.begin() + size_t(s + 1) * sz, d.memptr());
            out[s] = Rcpp::wrap(d);
        }
        return out;
    };
    Rcpp::List beta_draws_list = slices(beta_buf, 3, m, beta_sz);
    Rcpp::List threshold_draws_list = slices(thr_buf, m, C1, thr_sz);
    Rcpp::List f_draws_list = R_NilValue;
    Rcpp::List fstar_draws_list = R_NilValue;
    if (store_f) f_draws_list = slices(f_buf, n, m, f_sz);
    if (store_fstar) fstar_draws_list = slices(fstar_buf, GPIRT_NGRID, m, fstar_sz);

    return Rcpp::List::create(Rcpp::Named("theta", theta_draws), Rcpp::Named("f", f_draws_list), Rcpp::Named("beta", beta_draws_list),
                              Rcpp::Named("fstar", fstar_draws_list), Rcpp::Named("threshold", threshold_draws_list),
                              Rcpp::Named("ll", ll_draws));
}
