// TEST INFRASTRUCTURE — CPU oracle implementation (see orc_sampler.h for the reference citations).
#include "orc_sampler.h"

#include <chrono>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {

static inline double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static inline int thread_id() {
#ifdef _OPENMP
    return omp_get_thread_num();
#else
    return 0;
#endif
}
static inline int max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// Run body(rng) with the per-unit random source selected by s.
template <class F>
static inline bool with_rng_raw(const Src& s, size_t unit, size_t nz, F&& body) {
    if (s.mode == 0) {
        SuppliedRng r(s.z ? s.z + unit * s.zstride : nullptr, s.z ? nz : 0, s.u + unit * s.ustride, s.ustride);
        body(r);
        return r.overrun;
    } else if (s.mode == 1) {
        PhiloxRng r(s.seed, s.chain, s.iter, s.zs | (s.session << 8), s.us | (s.session << 8), uint32_t(unit));
        body(r);
        return false;
    } else {
        body((*s.mts)[thread_id()]);
        return false;
    }
}
template <class F>
static inline bool with_rng(const Src& s, size_t unit, size_t nz, F&& body) {
    if (!s.rec_z) return with_rng_raw(s, unit, nz, body);
    return with_rng_raw(s, unit, nz, [&](UnitRng& base) {
        RecordingRng r(base, (*s.rec_z)[unit], (*s.rec_u)[unit]);
        body(r);
    });
}

// The elliptical-slice transition shared by all four samplers (src/draw-f.cpp:14-41).
// log_y is supplied by the caller because the reference forms it in two different orders
// (ll + log u in draw-f/draw-beta; log u + terms in draw-theta/draw_threshold).
template <class Eval>
static inline void ess_loop(UnitRng& rng, double log_y, Eval&& eval_at, int32_t& evals, double& margin) {
    double eps_min = 0.0, eps_max = TWO_PI;
    double eps = eps_min + (eps_max - eps_min) * rng.unif();  // ThreadRNG::runif(min,max), src/gpirt.h:38-40
    eps_min = eps - TWO_PI;
    for (;;) {
        double llp = eval_at(std::cos(eps), std::sin(eps));
        ++evals;
        margin = std::min(margin, std::fabs(llp - log_y));
        if (llp > log_y) break;
        if (eps < 0.0) eps_min = eps; else eps_max = eps;
        eps = eps_min + (eps_max - eps_min) * rng.unif();
    }
}

void Sampler::init(size_t n_, size_t m_, size_t C_, const double* y_, const double* sds_, const double* tps_) {
    n = n_; m = m_; C = C_;
    y.assign(y_, y_ + n * m);
    sds.assign(sds_, sds_ + 3 * m);
    theta_prior_sds.assign(tps_, tps_ + 2 * n);
    obs_persons.assign(m, {});
    obs_items.assign(n, {});
    for (size_t j = 0; j < m; ++j)
        for (size_t i = 0; i < n; ++i)
            if (std::isfinite(y[i + j * n])) { obs_persons[j].push_back(i); obs_items[i].push_back(j); }
    theta.assign(n, 0.0); beta.assign(3 * m, 0.0); f.assign(n * m, 0.0); thr.assign(m * (C + 1), 0.0);
    fstar.assign(NGRID * m, 0.0); L.assign(n * n, 0.0);
    f_new.assign(n * m, 0.0); mu.assign(n * m, 0.0); mu_star.assign(NGRID * m, 0.0);
    mts.resize(max_threads());
    for (auto& d : diag) d.reset(0);
}

void Sampler::compute_mu() {
    // X = [1, theta, theta^2]; mu = X*beta; mu_star = Xstar*beta. Left-to-right accumulation.
    for (size_t j = 0; j < m; ++j) {
        const double b0 = beta[0 + 3 * j], b1 = beta[1 + 3 * j], b2 = (p_beta == 3) ? beta[2 + 3 * j] : 0.0;
        for (size_t i = 0; i < n; ++i) {
            double t = theta[i];
            double v = b0 + t * b1;
            if (p_beta == 3) v += (t * t) * b2;
            mu[i + j * n] = v;
        }
        for (size_t k = 0; k < NGRID; ++k) {
            double t = grid_value(int(k));
            double v = b0 + t * b1;
            if (p_beta == 3) v += (t * t) * b2;
            mu_star[k + j * NGRID] = v;
        }
    }
}

int Sampler::update_cholesky() {
    K(theta.data(), n, theta.data(), n, sds.data() /* column 0: item 0's sds */, L.data());
    for (size_t i = 0; i < n; ++i) L[i + i * n] += 1e-6;
    return chol_lower_blocked(L.data(), n);
}

void Sampler::draw_f(const Src& s, const double* nu_supplied) {
    double t0 = now();
    if (faithful) update_cholesky();  // src/draw-f.cpp:59 refactorises the same K
    t_fixed += now() - t0; t0 = now();
    diag[B_F].reset(m);
    bool over = false;
    BlasSerialScope blas_serial;
#pragma omp parallel for schedule(dynamic) reduction(|| : over)
    for (size_t j = 0; j < m; ++j) {
        std::vector<double> nu(n), z(n), fp(n), thr_j(C + 1);
        for (size_t c = 0; c <= C; ++c) thr_j[c] = thr[j + c * m];
        const double* fj = &f[j * n];
        const double* yj = &y[j * n];
        const double* muj = &mu[j * n];
        const auto& obs = obs_persons[j];
        bool o = with_rng(s, j, n, [&](UnitRng& rng) {
            if (nu_supplied) {
                std::memcpy(nu.data(), nu_supplied + j * n, n * sizeof(double));
            } else {
                for (size_t i = 0; i < n; ++i) z[i] = rng.norm();
                gemv(L.data(), n, n, z.data(), nu.data());  // dense, as src/mvnormal.h:23
            }
            double u = rng.unif();
            double log_y = ll_bar_sparse(fj, yj, muj, thr_j.data(), obs.data(), obs.size()) + std::log(u);
            diag[B_F].evals[j] = 1;
            ess_loop(rng, log_y, [&](double ce, double se) {
                for (size_t i = 0; i < n; ++i) fp[i] = fj[i] * ce + nu[i] * se;
                return ll_bar_sparse(fp.data(), yj, muj, thr_j.data(), obs.data(), obs.size());
            }, diag[B_F].evals[j], diag[B_F].margin[j]);
        });
        over = over || o;
        std::memcpy(&f_new[j * n], fp.data(), n * sizeof(double));
    }
    diag[B_F].overrun = over;
    f.swap(f_new);  // src/gpirtMCMC.cpp:274
    t_items += now() - t0;
}

int Sampler::draw_fstar(const Src& s) {
    const size_t N = NGRID;
    double t0 = now();
    std::vector<double> tstar(N);
    for (size_t k = 0; k < N; ++k) tstar[k] = grid_value(int(k));
    std::vector<double> kstar(n * N), tmp;
    K(theta.data(), n, tstar.data(), N, sds.data(), kstar.data());  // src/draw-fstar.cpp:26
    tmp = kstar;
    Lpost.assign(N * N, 0.0);
    if (blas().on) {  // timed baseline only (orc_blas.h): dtrtrs with N right-hand sides and the product tmp^T tmp as a BLAS-linked Armadillo runs them
        blas().dtrsm(102, 141, 122, 111, 131, int(n), int(N), 1.0, L.data(), int(n), tmp.data(), int(n));  // :30
        K(tstar.data(), N, tstar.data(), N, sds.data(), Lpost.data());                                     // :33
        blas().dsyrk(102, 122, 112, int(N), int(n), -1.0, tmp.data(), int(n), 1.0, Lpost.data(), int(N));   // :34 (lower triangle; chol reads only that)
    } else {
#pragma omp parallel for schedule(static)
        for (size_t k = 0; k < N; ++k) trsv_lower(L.data(), n, &tmp[k * n]);  // :30
        K(tstar.data(), N, tstar.data(), N, sds.data(), Lpost.data());  // :33
#pragma omp parallel for schedule(static)
        for (size_t b = 0; b < N; ++b) {
            for (size_t a = 0; a < N; ++a) {
                double acc = 0.0;
                const double* ta = &tmp[a * n];
                const double* tb = &tmp[b * n];
                for (size_t i = 0; i < n; ++i) acc += ta[i] * tb[i];
                Lpost[a + b * N] -= acc;  // :34
            }
        }
    }
    for (size_t k = 0; k < N; ++k) Lpost[k + k * N] += 1e-6;  // :35
    int info = chol_lower_blocked(Lpost.data(), N);            // :36
    t_fixed += now() - t0; t0 = now();
    if (info) return info;
    diag[B_FSTAR].reset(m);
    bool over = false;
    BlasSerialScope blas_serial;
#pragma omp parallel for schedule(dynamic) reduction(|| : over)
    for (size_t j = 0; j < m; ++j) {
        std::vector<double> alpha(&f[j * n], &f[j * n] + n), z(N), dz(N), mean(N);
        double_solve(L.data(), n, alpha.data());  // :51 (uses f, not f - mu)
        bool o = with_rng(s, j, N, [&](UnitRng& rng) {
            for (size_t k = 0; k < N; ++k) z[k] = rng.norm();
        });
        over = over || o;
        gemv(Lpost.data(), N, N, z.data(), dz.data());
        if (blas().on) blas().dgemv(102, 112, int(n), int(N), 1.0, kstar.data(), int(n), alpha.data(), 1, 0.0, mean.data(), 1);  // kstar^T alpha, :52
        for (size_t k = 0; k < N; ++k) {
            double acc = 0.0;
            if (blas().on) acc = mean[k];
            else {
                const double* ck = &kstar[k * n];
                for (size_t i = 0; i < n; ++i) acc += ck[i] * alpha[i];
            }
            fstar[k + j * N] = (acc + mu_star[k + j * N]) + dz[k];  // :52,55
        }
    }
    diag[B_FSTAR].overrun = over;
    t_items += now() - t0;
    return 0;
}

// arma::interp1(X, Y, XI, YI, "linear") as the reference uses it (src/draw-fstar.cpp:79-80): X sorted, duplicated abscissae dropped (the first
// in sorted order is kept; with std::stable_sort that is the lowest index — Armadillo's own sort is not stable, so WHICH duplicate it keeps is
// unspecified there), piecewise-linear in between, NaN outside [min X, max X].
static void interp1_linear(const double* X, const double* Y, size_t nx, const double* XI, size_t nq, double* YI) {
    std::vector<size_t> ord(nx);
    for (size_t i = 0; i < nx; ++i) ord[i] = i;
    std::stable_sort(ord.begin(), ord.end(), [&](size_t a, size_t b) { return X[a] < X[b]; });
    std::vector<double> xs, ys;
    for (size_t k : ord) {
        if (!xs.empty() && X[k] == xs.back()) continue;
        xs.push_back(X[k]);
        ys.push_back(Y[k]);
    }
    for (size_t q = 0; q < nq; ++q) {
        const double x = XI[q];
        if (xs.empty() || x < xs.front() || x > xs.back()) { YI[q] = std::numeric_limits<double>::quiet_NaN(); continue; }
        size_t hi = std::lower_bound(xs.begin(), xs.end(), x) - xs.begin();
        if (xs[hi] == x) { YI[q] = ys[hi]; continue; }
        const size_t lo = hi - 1;
        const double w = (x - xs[lo]) / (xs[hi] - xs[lo]);
        YI[q] = (1.0 - w) * ys[lo] + w * ys[hi];
    }
}

int Sampler::draw_fstar_cst(const Src& s) {
    const size_t N = NGRID, P = 100;  // n_induced_points, src/draw-fstar.cpp:73
    std::vector<double> tstar(N), tind(P);
    for (size_t k = 0; k < N; ++k) tstar[k] = grid_value(int(k));
    const double lo = *std::min_element(theta.begin(), theta.end()), hi = *std::max_element(theta.begin(), theta.end());
    const double delta = (hi - lo) / double(P - 1);  // arma::linspace: start + i * delta, last element = end
    for (size_t i = 0; i + 1 < P; ++i) tind[i] = lo + double(i) * delta;
    tind[P - 1] = hi;
    std::vector<double> find(P * m);
    for (size_t j = 0; j < m; ++j) interp1_linear(theta.data(), &f[j * n], n, tind.data(), P, &find[j * P]);  // :77-82
    std::vector<double> Lc(P * P);
    K(tind.data(), P, tind.data(), P, sds.data(), Lc.data());  // :85
    for (size_t k = 0; k < P; ++k) Lc[k + k * P] += 1e-6;
    int info = chol_lower_blocked(Lc.data(), P);
    if (info) return info;
    std::vector<double> kstar(P * N), tmp;
    K(tind.data(), P, tstar.data(), N, sds.data(), kstar.data());  // :90
    tmp = kstar;
    for (size_t k = 0; k < N; ++k) trsv_lower(Lc.data(), P, &tmp[k * P]);  // :92
    Lpost.assign(N * N, 0.0);
    K(tstar.data(), N, tstar.data(), N, sds.data(), Lpost.data());  // :93
#pragma omp parallel for schedule(static)
    for (size_t b = 0; b < N; ++b)
        for (size_t a = 0; a < N; ++a) {
            double acc = 0.0;
            for (size_t i = 0; i < P; ++i) acc += tmp[a * P + i] * tmp[b * P + i];
            Lpost[a + b * N] -= acc;  // :94
        }
    for (size_t k = 0; k < N; ++k) Lpost[k + k * N] += 1e-6;  // :95
    info = chol_lower_blocked(Lpost.data(), N);                // :96
    if (info) return info;
    diag[B_FSTAR].reset(m);
    bool over = false;
#pragma omp parallel for schedule(dynamic) reduction(|| : over)
    for (size_t j = 0; j < m; ++j) {
        std::vector<double> alpha(&find[j * P], &find[j * P] + P), z(N), dz(N);
        double_solve(Lc.data(), P, alpha.data());  // :113
        bool o = with_rng(s, j, N, [&](UnitRng& rng) {
            for (size_t k = 0; k < N; ++k) z[k] = rng.norm();
        });
        over = over || o;
        gemv(Lpost.data(), N, N, z.data(), dz.data());
        for (size_t k = 0; k < N; ++k) {
            double acc = 0.0;
            for (size_t i = 0; i < P; ++i) acc += kstar[k * P + i] * alpha[i];
            fstar[k + j * N] = (acc + mu_star[k + j * N]) + dz[k];  // :114,117 (mu_star of session 0)
        }
    }
    diag[B_FSTAR].overrun = over;
    return 0;
}

// compute_ll_sparse (src/draw-theta.cpp:4-24) for respondent i at a raw theta value
static inline double theta_ll(const Sampler& S, size_t i, double theta, const double* fst, const double* must, const double* thr = nullptr) {
    const size_t N = NGRID;
    if (!thr) thr = S.thr.data();
    int idx = grid_index(theta);
    double result = 0.0;
    for (uint64_t j : S.obs_items[i]) {
        int c = static_cast<int>(S.y[i + j * S.n]);
        double g = fst[idx + j * N] + must[idx + j * N];
        result += cell_ll(g, c, &thr[j], S.m);
    }
    return result;
}

void Sampler::draw_theta(const Src& s) {
    diag[B_THETA].reset(n);
    std::vector<double> theta_new(n);
    bool over = false;
#pragma omp parallel for schedule(dynamic) reduction(|| : over)
    for (size_t i = 0; i < n; ++i) {
        const double* fst = fstar.data();
        const double* must = mu_star.data();
        std::vector<double> fcopy, mcopy;
        if (faithful) {  // src/draw-theta.cpp:144-156 copies both N x m tables row by row, per respondent
            fcopy.resize(NGRID * m); mcopy.resize(NGRID * m);
            for (size_t k = 0; k < NGRID; ++k)
                for (size_t j = 0; j < m; ++j) {
                    fcopy[k + j * NGRID] = fstar[k + j * NGRID];
                    mcopy[k + j * NGRID] = mu_star[k + j * NGRID];
                }
            fst = fcopy.data(); must = mcopy.data();
        }
        double Li = theta_regime == 0 ? std::sqrt(1.0 + std::pow(theta_prior_sds[0 + 2 * i], 2))  // :158
                                      : std::sqrt(theta_os * theta_os + 1e-6);                     // cholesky-cache.cpp:30-33, H=1
        double raw = 0.0;
        bool o = with_rng(s, i, 1, [&](UnitRng& rng) {
            double nu = Li * rng.norm();
            double u = rng.unif();
            double log_y = std::log(u);
            log_y += theta_ll(*this, i, theta[i], fst, must);
            diag[B_THETA].evals[i] = 1;
            ess_loop(rng, log_y, [&](double ce, double se) {
                raw = theta[i] * ce + nu * se;
                raw = std::max(-5.0, std::min(5.0, raw));
                double lp = 0.0;
                lp += theta_ll(*this, i, raw, fst, must);
                return lp;
            }, diag[B_THETA].evals[i], diag[B_THETA].margin[i]);
        });
        over = over || o;
        theta_new[i] = grid_value(grid_index(raw));  // :165-169
    }
    diag[B_THETA].overrun = over;
    theta.swap(theta_new);
}

void panel_draw_theta(Sampler* const* S, size_t H, int regime, const double* L_time, const Src* srcs) {
    const size_t n = S[0]->n;
    std::vector<std::vector<double>> theta_new(H, std::vector<double>(n));
    if (regime == 1) {
        for (size_t h = 0; h < H; ++h) S[h]->diag[B_THETA].reset(n);
        std::vector<char> over(H, 0);
#pragma omp parallel for schedule(dynamic)
        for (size_t i = 0; i < n; ++i) {
            for (size_t h = 0; h < H; ++h) {  // src/draw-theta.cpp:182-202
                Sampler& Sh = *S[h];
                // QUIRK kept: the RDM branch hands the whole thresholds cube to the 1-session ESS (src/draw-theta.cpp:196-198), which reads
                // thresholds.slice(0) (:44-47) — every session's theta is scored against SESSION 0's thresholds.
                const double* thr0 = S[0]->thr.data();
                const double Li = std::sqrt(1.0 + std::pow(Sh.theta_prior_sds[0 + 2 * i], 2));  // :194
                double raw = 0.0;
                Diag& d = Sh.diag[B_THETA];
                bool o = with_rng(srcs[h], i, 1, [&](UnitRng& rng) {
                    double nu = Li * rng.norm();
                    double u = rng.unif();
                    double log_y = std::log(u);
                    log_y += theta_ll(Sh, i, Sh.theta[i], Sh.fstar.data(), Sh.mu_star.data(), thr0);
                    d.evals[i] = 1;
                    ess_loop(rng, log_y, [&](double ce, double se) {
                        raw = Sh.theta[i] * ce + nu * se;
                        raw = std::max(-5.0, std::min(5.0, raw));
                        double lp = 0.0;
                        lp += theta_ll(Sh, i, raw, Sh.fstar.data(), Sh.mu_star.data(), thr0);
                        return lp;
                    }, d.evals[i], d.margin[i]);
                });
                if (o) {
#pragma omp critical
                    over[h] = 1;
                }
                theta_new[h][i] = grid_value(grid_index(raw));  // :199-201
            }
        }
        for (size_t h = 0; h < H; ++h) S[h]->diag[B_THETA].overrun = over[h] != 0;
    } else {
        Diag& d = S[0]->diag[B_THETA];
        d.reset(n);
        bool over = false;
#pragma omp parallel for schedule(dynamic) reduction(|| : over)
        for (size_t i = 0; i < n; ++i) {
            std::vector<double> z(H), nu(H), raw(H, 0.0);
            bool o = with_rng(srcs[0], i, H, [&](UnitRng& rng) {
                for (size_t h = 0; h < H; ++h) z[h] = rng.norm();
                gemv(L_time, H, H, z.data(), nu.data());  // rmvnorm_threadsafe(L, rng): dense L * z, src/mvnormal.h:17-24
                double u = rng.unif();
                double log_y = std::log(u);
                for (size_t h = 0; h < H; ++h) log_y += theta_ll(*S[h], i, S[h]->theta[i], S[h]->fstar.data(), S[h]->mu_star.data());  // :43-47
                d.evals[i] = 1;
                ess_loop(rng, log_y, [&](double ce, double se) {
                    for (size_t h = 0; h < H; ++h) {
                        raw[h] = S[h]->theta[i] * ce + nu[h] * se;
                        raw[h] = std::max(-5.0, std::min(5.0, raw[h]));  // :58-62
                    }
                    double lp = 0.0;
                    for (size_t h = 0; h < H; ++h) lp += theta_ll(*S[h], i, raw[h], S[h]->fstar.data(), S[h]->mu_star.data());
                    return lp;
                }, d.evals[i], d.margin[i]);
            });
            over = over || o;
            for (size_t h = 0; h < H; ++h) theta_new[h][i] = grid_value(grid_index(raw[h]));  // :229-233
        }
        d.overrun = over;
    }
    for (size_t h = 0; h < H; ++h) S[h]->theta.swap(theta_new[h]);
}

void Sampler::gather_f() {  // src/gpirtMCMC.cpp:291-298
    for (size_t k = 0; k < n; ++k) {
        int idx = grid_index(theta[k]);
        for (size_t j = 0; j < m; ++j) f[k + j * n] = fstar[idx + j * NGRID];
    }
}

void Sampler::draw_beta(const Src& s) {
    diag[B_BETA].reset(m);
    std::vector<double> beta_new(beta);
    const int p = p_beta;
    bool over = false;
#pragma omp parallel for schedule(dynamic) reduction(|| : over)
    for (size_t j = 0; j < m; ++j) {
        const auto& obs = obs_persons[j];
        if (obs.empty()) continue;  // src/draw-beta.cpp:73,97-99: beta kept, no random numbers drawn
        const size_t no = obs.size();
        // extract observed rows (src/draw-beta.cpp:75-79)
        std::vector<double> f_obs(no), y_obs(no), X1(no), X2(no), mu_obs(no), thr_j(C + 1);
        std::vector<uint64_t> valid(no);
        for (size_t k = 0; k < no; ++k) {
            size_t i = obs[k];
            f_obs[k] = f[i + j * n]; y_obs[k] = y[i + j * n];
            X1[k] = theta[i]; X2[k] = theta[i] * theta[i];  // arma::pow(theta,2), src/gpirtMCMC.cpp:288
            valid[k] = k;
        }
        for (size_t c = 0; c <= C; ++c) thr_j[c] = thr[j + c * m];
        double cholS[3], b[3] = {0, 0, 0}, bp[3] = {0, 0, 0}, nu[3] = {0, 0, 0};
        for (int q = 0; q < 3; ++q) cholS[q] = std::sqrt(sds[q + 3 * j] * sds[q + 3 * j] + 1e-6);  // :82-86
        for (int q = 0; q < p; ++q) b[q] = beta[q + 3 * j];
        auto xb = [&](const double* bb) {
            for (size_t k = 0; k < no; ++k) {
                double v = bb[0] + X1[k] * bb[1];
                if (p == 3) v += X2[k] * bb[2];
                mu_obs[k] = v;
            }
        };
        bool o = with_rng(s, j, 3, [&](UnitRng& rng) {
            // rmvnorm_threadsafe(cholS_small): always 3 normals (cholS_small is 3x3, src/gpirt.h:128)
            double z[3];
            for (int q = 0; q < 3; ++q) z[q] = rng.norm();
            for (int q = 0; q < 3; ++q) nu[q] = cholS[q] * z[q];
            double u = rng.unif();
            xb(b);
            double log_y = ll_bar_sparse(f_obs.data(), y_obs.data(), mu_obs.data(), thr_j.data(), valid.data(), no) + std::log(u);
            diag[B_BETA].evals[j] = 1;
            ess_loop(rng, log_y, [&](double ce, double se) {
                for (int q = 0; q < p; ++q) bp[q] = b[q] * ce + nu[q] * se;
                xb(bp);
                return ll_bar_sparse(f_obs.data(), y_obs.data(), mu_obs.data(), thr_j.data(), valid.data(), no);
            }, diag[B_BETA].evals[j], diag[B_BETA].margin[j]);
        });
        over = over || o;
        for (int q = 0; q < p; ++q) beta_new[q + 3 * j] = bp[q];
    }
    diag[B_BETA].overrun = over;
    beta.swap(beta_new);
}

void Sampler::draw_threshold(const Src& s) {
    diag[B_THR].reset(m);
    std::vector<double> thr_new(thr);
    bool over = false;
#pragma omp parallel for schedule(dynamic) reduction(|| : over)
    for (size_t j = 0; j < m; ++j) {
        const auto& obs = obs_persons[j];
        std::vector<double> thr_j(C + 1), delta(C - 1), dp(C - 1), nu(C - 1), tp(C + 1);
        for (size_t c = 0; c <= C; ++c) thr_j[c] = thr[j + c * m];
        threshold_to_delta(thr_j.data(), C, delta.data());  // src/draw_threshold.cpp:219
        const double* fj = &f[j * n];
        const double* yj = &y[j * n];
        const double* muj = &mu[j * n];
        auto ll_with = [&](const double* t, double start) {
            double acc = start;
            for (uint64_t i : obs) acc += cell_ll(fj[i] + muj[i], static_cast<int>(yj[i]), t);
            return acc;
        };
        bool o = with_rng(s, j, C - 1, [&](UnitRng& rng) {
            for (size_t q = 0; q + 1 < C; ++q) nu[q] = rng.norm();  // identity prior covariance, :112-114
            double u = rng.unif();
            delta_to_threshold(delta.data(), C, tp.data());  // :118
            double log_y = ll_with(tp.data(), std::log(u));
            diag[B_THR].evals[j] = 1;
            ess_loop(rng, log_y, [&](double ce, double se) {
                for (size_t q = 0; q + 1 < C; ++q) dp[q] = delta[q] * ce + nu[q] * se;
                delta_to_threshold(dp.data(), C, tp.data());
                return ll_with(tp.data(), 0.0);
            }, diag[B_THR].evals[j], diag[B_THR].margin[j]);
        });
        over = over || o;
        delta_to_threshold(dp.data(), C, tp.data());  // :226
        for (size_t c = 0; c <= C; ++c) thr_new[j + c * m] = tp[c];
    }
    diag[B_THR].overrun = over;
    thr.swap(thr_new);
}

double Sampler::total_ll() {  // src/gpirtMCMC.cpp:324-331
    std::vector<double> per(m);
#pragma omp parallel for schedule(static)
    for (size_t j = 0; j < m; ++j) {
        std::vector<double> thr_j(C + 1);
        for (size_t c = 0; c <= C; ++c) thr_j[c] = thr[j + c * m];
        per[j] = ll_bar_sparse(&f[j * n], &y[j * n], &mu[j * n], thr_j.data(), obs_persons[j].data(), obs_persons[j].size());
    }
    double acc = 0.0;
    for (size_t j = 0; j < m; ++j) acc += per[j];
    return acc;
}

// One sweep, src/gpirtMCMC.cpp:267-331. Returns 0, or the failing minor of a Cholesky (+1000000 for K_post).
int Sampler::sweep(const Src srcs[N_BLOCKS]) {
    draw_f(srcs[B_F], nullptr);
    int info = draw_fstar(srcs[B_FSTAR]);
    if (info) return 1000000 + info;
    double t0 = now();
    draw_theta(srcs[B_THETA]);
    gather_f();
    draw_beta(srcs[B_BETA]);
    compute_mu();
    t_items += now() - t0; t0 = now();
    info = update_cholesky();
    t_fixed += now() - t0; t0 = now();
    if (info) return info;
    draw_threshold(srcs[B_THR]);
    ll = total_ll();
    t_items += now() - t0;
    return 0;
}

}  // namespace orc
