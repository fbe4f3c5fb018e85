// The sampler handle behind the C-ABI: owns every device buffer of one chain's item shard.
// Replaces the reference's CholeskyCache + WorkspacePool + the state variables of gpirtMCMC()
// (src/gpirt.h:49-68,153-195; src/gpirtMCMC.cpp:89-90,138-224).
//
// HBM layout (all fp64 matrices column-major, leading dimensions padded to 16 doubles, padding zero):
//   y_im   int8  [m_local][ldy]            item-major responses of the local items (0 = missing)
//   y_rm   int8  [n_local][ldm]            respondent-major rows of the local respondents, all items
//   theta  f64   [n]                       current abilities (always on the theta* grid after sweep 1)
//   theta_x f64  [world*n_chunk]           all-gather buffer of the theta step (chunk r = respondents of rank r)
//   beta   f64   [3][world*m_chunk]        all-gather buffer, chunk r = items of rank r
//   thr    f64   [C+1][world*m_chunk]      item-major thresholds (transposed w.r.t. the reference's m x (C+1))
//   f, nu  f64   [ldn][m_local]            latent responses and the prior draws L*Z
//   zbuf   f64   [max(ldn,ldN)][m_local]   normals: Z (n x m) for draw_f, then Z* (N x m) for draw_fstar
//   fstar  f64   [world][NGRID][ld_chunk]  grid-row-major f* table: the row of one grid point is contiguous
//   L      f64   [ldn][n]                  chol(K(theta,theta) + 1e-6 I)
//   W      f64   [ldn][NGRID]              k*, then L^-1 k*, then K^-1 k*
//   Kpost  f64   [ldN][NGRID]              K** - tmp^T tmp + 1e-6 I, then its Cholesky factor
//   Linv   f64   [ldn][n]                  L^-1 (solver 1), W2 [ldn][NGRID] = L^-1 k*
#pragma once
#include <cuda_runtime.h>

#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "../../include/gpirt_b200.h"
#include "linalg.cuh"
#include "samplers.cuh"

struct gpirt_sampler {
    gpirt_config cfg{};
    int n = 0, m_total = 0, C = 0, world = 1, rank = 0;
    int m_chunk = 0, m_local = 0, item0 = 0;
    int n_chunk = 0, n_local = 0, resp0 = 0;
    int ldn = 0, ldN = 0, ldy = 0, ld_chunk = 0, ldz = 0;
    long long ldm = 0;
    int theta_regime = 0;

    cudaStream_t stream = nullptr;
    bool own_stream = false;
    uint32_t iter = 0;
    bool have_data = false, chol_valid = false;
    // The theta-only chain of a sweep (K, chol(K), k*, L^-1, K_post, chol(K_post), K^-1 k*) runs on a second, higher-priority
    // stream as soon as the new theta is committed, under the item kernels of the main stream (draw_beta, draw_threshold,
    // ll, next draw_f): it is a long chain of small launches that cannot fill the GPU on its own.
    cudaStream_t side = nullptr;
    cudaStream_t side2 = nullptr;    // W = L^-T tmp runs here, beside K_post and its Cholesky on `side` (both need tmp only)
    cudaEvent_t ev_tmp = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_L = nullptr, ev_W = nullptr, ev_shared = nullptr;
    cudaEvent_t ev_xchg = nullptr;  // "this shard's chunk is written": what gpirt_exchange_pull makes the pulling shard's stream wait for
    gpirt::CholMilestones L_rows;  // leading row blocks of L signalled while chol(K) is still running on the side stream (nu = L Z starts under it)
    bool pending_rows = false;
    gpirt::CholAux chol_aux;       // helper stream + events of the two-level Cholesky's look-ahead (n >= 4096)
    gpirt::CholMilestones Lpost_rows;  // the same for chol(K_post): the f* noise term Z*^T L_post^T is launched in column slices under it
    bool pending_post_rows = false;
    bool overlap = true;           // gpirt_set_overlap; profiling forces the serial order
    bool shared_valid = false;     // W / Kpost hold the draw_fstar terms of the current theta
    bool pending_L = false, pending_W = false, pending_shared = false;  // the main stream has not yet waited for ev_L / ev_W / ev_shared

    int8_t *y_im = nullptr, *y_rm = nullptr;
    // item-sharded theta step (world > 1, samplers.cuh ThetaItemsView): all respondents' rows over the LOCAL items, the partial-sum exchange buffer,
    // per-respondent round state, and the packed (beta, thresholds, item ll) exchange buffer
    int8_t* y_rl = nullptr;
    long long ldml = 0;
    double *theta_part = nullptr, *theta_ll0 = nullptr, *items_packed = nullptr;
    int32_t *theta_knext = nullptr, *theta_nact = nullptr;
    const double* theta_z_dev = nullptr; long long theta_z_stride = 0;  // supplied streams of the current item-sharded theta step (staged at begin)
    gpirt::UniformSrc theta_us{nullptr, 0};
    double *theta = nullptr, *theta_x = nullptr, *beta = nullptr, *thr = nullptr;
    double *f = nullptr, *nu = nullptr, *zbuf = nullptr, *fstar = nullptr;
    double *L = nullptr, *W = nullptr, *Kpost = nullptr, *tstar = nullptr;
    double *Linv = nullptr, *W2 = nullptr, *tri_scratch = nullptr;  // block-inverse solver: L^-1, L^-1 k*, merge scratch (allocated on first use)
    int solver = 1;  // 0: blocked substitution TRSM (backward stable), 1: explicit L^-1 by parallel block inversion + GEMMs
    double* mu_rows = nullptr;    // optional per-row means of the f / threshold blocks (gpirt_set_mean_rows), [ldn][m_local]
    double* theta_tab = nullptr;  // [NGRID][m_total][C] per-sweep table of theta-step terms (allocated on first use)
    size_t theta_tab_bytes = 0;
    double *sds = nullptr, *tps = nullptr, *item_ll = nullptr, *total_ll = nullptr;
    int32_t* evals[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    int32_t* exact[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // per unit: evaluations that ran in fp64 (screening diagnostics)
    int screen = 1;  // fp32 screening of ESS proposals with a rigorous error bound; decisions identical to fp64 (gpirt_set_screen)
    // screened proposals stop at the first partial sum that already decides a rejection. OFF by default: measured on B200 at C3 (profiles/
    // r02_ess_early_exit.txt) only 4 % / 22 % of the proposals are decided after 1/8 / 3/8 of the cells (the cap on the unseen terms is 0, so a proposal
    // must be 8x / 2.7x worse than the state) and each extra block reduction costs ~14 % of a pass: 0.63 -> 0.68 ms for the f block.
    int early_exit = getenv("GPIRT_ESS_EARLY") ? atoi(getenv("GPIRT_ESS_EARLY")) != 0 : 0;
    double* margin[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    gpirt::DevFlags* flags = nullptr;
    int* chol_ws = nullptr;  // per-panel arrival counters of the blocked Cholesky
    // persistent single-kernel Cholesky (chol_persistent.cu): task lists of the n x n and the NGRID x NGRID factorisation, one state block each
    uint32_t *chol_tasks_n = nullptr, *chol_tasks_N = nullptr;
    int *chol_state_n = nullptr, *chol_state_N = nullptr;
    gpirt::CholPlan plan_n, plan_N;
    double sds0[3] = {0, 0, 0};  // beta_prior_sds.col(0): the sds every K() call uses (src/cholesky-cache.cpp:21)

    // optional per-section device timing (gpirt_profile_*): CUDA events around each kernel group
    struct Pending { const char* name; cudaEvent_t a, b; };
    bool profiling = false;
    int profile_level = 0;  // 1: serial per-kernel times; 2: timeline of the overlapped schedule (times include waits / SM sharing)
    std::vector<Pending> pending;
    std::map<std::string, std::pair<double, long>> sections;  // name -> (total ms, launches)

    // staging for supplied streams
    double *d_z = nullptr, *d_u = nullptr;
    int session = 0;                    // horizon > 1: which session of a gpirt_panel this handle is (keys the Philox streams)
    const double* thr_theta = nullptr;  // thresholds the theta step scores against when they are not this handle's own (RDM quirk, capi.cu)
    char* arena = nullptr;  // the one allocation behind every buffer above
    // f* exchange through NVLink peer memory (item-sharded sweeps): peers' arenas opened with cudaIpc, addressed at the same offsets as ours
    unsigned long long* flags_x = nullptr;      // [MAX_PEERS + 1] arrival epochs, word p written by rank p
    char* peer_arena[gpirt::MAX_PEERS + 1] = {};  // index = peer rank
    size_t peer_fstar_off[gpirt::MAX_PEERS + 1] = {}, peer_flags_off[gpirt::MAX_PEERS + 1] = {};  // where the peer keeps its f* table / flag words
    int peers_attached = 0;
    unsigned long long push_epoch = 0;
    // Distributed theta-only chain (item shards with n >= 4096, capi.cu "distributed chain"): the peers' L / L^-1 / W / W2 matrices and chain flag words.
    enum { OFF_L = 0, OFF_LINV = 1, OFF_W = 2, OFF_W2 = 3, OFF_CFLAGS = 4, N_CHAIN_OFFS = 5 };
    unsigned long long* flags_c = nullptr;             // [MAX_PEERS + 1] chain epochs, word p written by rank p (monotonic)
    char* chain_peer[gpirt::MAX_PEERS + 1] = {};       // base of peer p's arena as this process / device can address it (IPC mapping or the pointer itself)
    bool chain_peer_ipc[gpirt::MAX_PEERS + 1] = {};    // opened here with cudaIpcOpenMemHandle (closed at detach)
    size_t chain_off[gpirt::MAX_PEERS + 1][N_CHAIN_OFFS] = {};
    int chain_attached = 0;
    unsigned long long chain_epoch = 0;                // advanced identically on every rank (the exchange points of the chain are a fixed program order)
    bool dist_chain = false;
    bool kstar_built = false;                          // k* is in W already (the distributed chain builds it before its start barrier)
    size_t d_z_cap = 0, d_u_cap = 0;
    // device-side store of the draws gpirtMCMC keeps (src/gpirtMCMC.cpp:230-255,334-356): appended on the stream, read back in bulk
    struct DrawStore {
        int cap = 0, count = 0;
        double *theta = nullptr, *beta = nullptr, *thr = nullptr, *ll = nullptr, *f = nullptr, *fstar = nullptr;  // [cap][per-sample block]
    } store;

    gpirt::RngKey key() const {
        gpirt::RngKey k;
        k.k0 = (uint32_t)cfg.seed;
        k.k1 = (uint32_t)(cfg.seed >> 32);
        k.tag = ((uint32_t)session << 8) | (cfg.chain << 16);
        k.iter = iter;
        return k;
    }
};
