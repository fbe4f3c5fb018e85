// Drop-in replacement for the body of recover_fstar() (reference: src/recover_fstar.cpp:8-93) over libgpirt_b200.so: one draw_f and one
// draw_fstar from a stored f under a given seed; returns list(fstar = NGRID x m x horizon), as src/recover_fstar.cpp:92.
#include <RcppArmadillo.h>

#include "gpirt_b200.h"

// [[Rcpp::export(.recover_fstar)]]
Rcpp::List recover_fstar(int seed, arma::cube f, const arma::cube& y, const arma::mat& theta, const arma::cube& beta,
                         const arma::cube& thresholds, const arma::mat& beta_prior_means, const arma::mat& beta_prior_sds,
                         const int constant_IRF) {
    arma::cube fstar(GPIRT_NGRID, f.n_cols, f.n_slices);
    gpirt_recover_fstar_args a{};
    a.seed = seed; a.f = f.memptr(); a.y = y.memptr(); a.n = int(f.n_rows); a.m = int(f.n_cols); a.horizon = int(f.n_slices);
    a.theta = theta.memptr(); a.beta = beta.memptr(); a.thresholds = thresholds.memptr(); a.C = int(thresholds.n_cols) - 1;
    a.beta_prior_means = beta_prior_means.memptr(); a.beta_prior_sds = beta_prior_sds.memptr();
    a.constant_IRF = constant_IRF; a.device = 0; a.fstar = fstar.memptr();
    if (gpirt_recover_fstar(&a) != GPIRT_OK) Rcpp::stop(gpirt_last_error());
    return Rcpp::List::create(Rcpp::Named("fstar", fstar));
}
