// TEST INFRASTRUCTURE — CPU oracle: the per-block samplers and the sweep (see orc_math.h header).
// Restated from (citations relative to /root/reference):
//   ESS skeleton              src/draw-f.cpp:5-45 (identical in draw-theta.cpp:26-84, draw-beta.cpp:4-49,
//                             draw_threshold.cpp:99-169)
//   draw_f                    src/draw-f.cpp:47-83           (constant_IRF == 0, horizon == 1)
//   draw_fstar                src/draw-fstar.cpp:18-57
//   draw_theta                src/draw-theta.cpp:4-24,112-170 (+ the L_i of :171-236 at horizon 1)
//   f <- f* gather            src/gpirtMCMC.cpp:291-298
//   draw_beta                 src/draw-beta.cpp:51-101
//   draw_threshold            src/draw_threshold.cpp:206-228
//   sweep order / total ll    src/gpirtMCMC.cpp:261-331
//   constant_IRF == 1         src/draw-fstar.cpp:58-125 (Sampler::draw_fstar_cst); the stacked-session ESS of src/draw-f.cpp:84-139 and
//                             src/draw_threshold.cpp:5-96,181-205 are the single-session blocks run on a sampler over the n*H stacked rows
//                             (oracle/orc.py ConstantPanel)
// Scope: constant_IRF == 0 in this file's sweep; horizon == 1 (every BASELINE.json config) is one Sampler; horizon > 1 is a panel of H Samplers (sessions are
// independent given theta: src/draw-f.cpp:62, src/draw-fstar.cpp:20, src/draw-beta.cpp:60, src/draw_threshold.cpp:206) plus
// panel_draw_theta below for the RDM and GP-over-time regimes; the CST regime at horizon > 1 indexes thresholds out of bounds in the
// reference (src/draw-theta.cpp:162 passes the m-row cube for m*H stacked items) and is not restated. Parity of this file itself
// is pinned against the reference's own translation units compiled under oracle/_ref (see
// oracle/Makefile and tests/test_oracle_vs_ref.py).
#pragma once
#include <algorithm>
#include <cstring>
#include <vector>

#include "orc_math.h"
#include "orc_rng.h"

namespace orc {

static const double TWO_PI = 6.283185307179586476925286766559;  // Rmath M_2PI
static const size_t NGRID = 1001;                               // regspace(-5, 0.01, 5)

inline double grid_value(int idx) { return -5.0 + double(idx) * 0.01; }
inline int grid_index(double theta) {
    int idx = static_cast<int>(std::round((theta + 5.0) / 0.01));
    return std::max(0, std::min(idx, int(NGRID) - 1));
}

enum Block { B_F = 0, B_FSTAR = 1, B_THETA = 2, B_BETA = 3, B_THR = 4, N_BLOCKS = 5 };

// Where random numbers come from for one call of a block sampler.
struct Src {
    int mode = 0;  // 0 supplied, 1 philox, 2 reference ThreadRNG (mt19937)
    // supplied: z[unit*zstride + i], u[unit*ustride + k]
    const double* z = nullptr; size_t zstride = 0;
    const double* u = nullptr; size_t ustride = 0;
    // philox
    uint64_t seed = 0; uint32_t chain = 0, iter = 0, zs = 0, us = 0;
    uint32_t session = 0;  // horizon > 1: folded into the stream id (stream | session << 8), as gpirt_b200/csrc/rng.cuh does
    // mt: one generator per worker thread
    std::vector<MtRng>* mts = nullptr;
    // when set, every unit's draws are appended to (*rec_z)[unit] / (*rec_u)[unit]
    std::vector<std::vector<double>>* rec_z = nullptr;
    std::vector<std::vector<double>>* rec_u = nullptr;
};

struct Diag {  // per-unit diagnostics of one block call
    std::vector<int32_t> evals;    // likelihood evaluations (1 + proposals)
    std::vector<double> margin;    // min |ll' - log_y| over the unit's proposals
    bool overrun = false;          // a supplied stream ran dry
    void reset(size_t n) { evals.assign(n, 0); margin.assign(n, std::numeric_limits<double>::infinity()); overrun = false; }
};

struct Sampler {
    size_t n = 0, m = 0, C = 0;
    int p_beta = 3;            // rows of beta: 3 in gpirtMCMC, 2 in recover_fstar (src/recover_fstar.cpp:58-60)
    int faithful = 0;          // 1: keep the reference's redundant work (second chol, table copies)
    int theta_regime = 0;      // 0 CST / RDM: L_i = sqrt(1+sd0_i^2); 1 GP-over-time at H=1: sqrt(os^2+1e-6)
    double theta_os = 1.0;
    std::vector<double> y;                 // n x m, NaN missing
    std::vector<std::vector<uint64_t>> obs_persons, obs_items;
    std::vector<double> sds;               // beta_prior_sds 3 x m
    std::vector<double> theta_prior_sds;   // 2 x n
    // state
    std::vector<double> theta, beta, f, thr, fstar, L;
    double ll = 0.0;
    // wall-clock split of the sweeps run so far: work shared by all items (Cholesky of K, k*, K_post) vs per-item /
    // per-respondent work, which is linear in m (used by bench.py to extrapolate a bounded sample to the full item count)
    double t_fixed = 0.0, t_items = 0.0;
    // scratch
    std::vector<double> f_new, mu, mu_star, Lpost, W;
    std::vector<MtRng> mts;
    std::vector<MtRng>* mts_shared = nullptr;  // horizon > 1: the sessions of a panel consume ONE set of ThreadRNGs, as the reference's ws_pool
    Diag diag[N_BLOCKS];
    bool record = false;  // keep the draws of the last call of each block (per unit)
    std::vector<std::vector<double>> rec_z[N_BLOCKS], rec_u[N_BLOCKS];

    void init(size_t n_, size_t m_, size_t C_, const double* y_, const double* sds_, const double* tps_);
    void compute_mu();        // mu = X beta, mu_star = X* beta (src/gpirtMCMC.cpp:308-311)
    int update_cholesky();    // src/cholesky-cache.cpp:20-24
    void draw_f(const Src& s, const double* nu_supplied);
    int draw_fstar(const Src& s);
    // constant_IRF == 1 (src/draw-fstar.cpp:58-125) on the STACKED sampler of a constant-IRF panel (theta, f: the n*H stacked rows; beta:
    // session 0's): 100 inducing points on linspace(min theta, max theta), f interpolated onto them (arma::interp1, "linear": abscissae sorted,
    // duplicates dropped keeping the first), GP posterior on the grid from the inducing values, one f* for all sessions.
    int draw_fstar_cst(const Src& s);
    void draw_theta(const Src& s);
    void gather_f();
    void draw_beta(const Src& s);
    void draw_threshold(const Src& s);
    double total_ll();
    int sweep(const Src srcs[N_BLOCKS]);
};

// draw_theta at horizon > 1 (src/draw-theta.cpp:171-236) over the H single-session samplers of a panel.
//   regime 1 (RDM, ls <= 0.1):           every (respondent, session) is its own 1-dim ESS, L_i = sqrt(1 + sd0_i^2); loop order i, then h (:178-203);
//                                        srcs[h] feeds session h (unit = respondent)
//   regime 2 (GP over time, 0.1<ls<3H):  one H-dim ESS per respondent with nu = L_time z (:206-235); srcs[0] feeds it (H normals per unit)
// Diagnostics / recordings land in each session's diag[B_THETA] (RDM) or in session 0's (GP).
void panel_draw_theta(Sampler* const* S, size_t H, int regime, const double* L_time, const Src* srcs);

}  // namespace orc
