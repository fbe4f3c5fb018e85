// Device-side views and launchers of the per-block samplers (see samplers.cu).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "rng.cuh"

namespace gpirt {

extern long long g_kernel_launches;  // defined in linalg.cu

constexpr int NGRID = 1001;
constexpr int MAX_C = 21;
constexpr int ESS_MAX_PROPOSALS = 2000;  // the reference has no cap (src/draw-f.cpp:23); a NaN likelihood would spin forever

// flags raised by kernels, read back at sync points
struct DevFlags {
    int chol_info;       // failing leading minor of chol(K)
    int chol_post_info;  // failing leading minor of chol(K_post)
    int overrun;         // a supplied uniform stream ran dry
    int stalled;         // an ESS loop hit ESS_MAX_PROPOSALS
    int bad_code;        // a y code outside 0..C
    int peer_timeout;    // a peer's f* chunk or chain block did not arrive within the wait limit (item-sharded sweeps over NVLink peer memory)
};

// Item-sharded sweeps: the f* exchange that feeds the theta step (the all-gather of SURVEY.md 8e) is done by the producing kernel itself.
// Rank r's k_add_mustar_push finishes its chunk of the f* table and stores every finished value into chunk r of every peer's table through
// NVLink peer mappings (cudaIpc), k_signal_peers publishes the epoch in the peers' flag words, k_wait_peers holds the theta step back
// until all chunks of this epoch have landed.
constexpr int MAX_PEERS = 15;
struct PeerPush {
    int count;                              // world - 1
    double* chunk[MAX_PEERS];               // peer p's address of chunk `rank` of its f* table
    unsigned long long* flag[MAX_PEERS];    // peer p's flag word for `rank`
};

// Everything the item-parallel kernels need (one warp per local item).
struct ItemView {
    int n, ldn;               // respondents; leading dimension of f / nu
    int m_local, item0;       // local items and the global index of the first one
    int m_total, C, p_beta;
    const int8_t* y_im;       // item-major responses of the local items: y_im[j*ldy + i], 0 = missing
    int ldy;
    const double* theta;      // n
    double* beta;             // 3 x m_total
    double* thr;              // (C+1) x m_total item-major: thr[J*(C+1) + c]
    double* f;                // n x m_local
    const double* nu;         // n x m_local (draw_f)
    const double* sds;        // 3 x m_total beta_prior_sds
    const double* mu_rows;    // optional n x m_local means (leading dimension ldn) used INSTEAD of X beta in the f and threshold blocks: the
                              // stacked sampler of a constant-IRF panel, whose rows carry their own session's X_h beta_h (src/draw-f.cpp:93-104)
    const double* z; long long z_stride;   // supplied normals (beta / thresholds) or nullptr
    UniformSrc us;
    RngKey key;
    int32_t* evals;           // per global item
    int32_t* exact_evals;     // per global item: evaluations that needed the fp64 pass (screening diagnostics), or nullptr
    int screen;               // 1: fp32 screening of proposals with a rigorous error bound (probit.cuh)
    int early_exit;           // 1 (with screen): a proposal is rejected as soon as a partial sum plus the cap of the remaining terms is below log_y
    double* margin;
    double* item_ll;          // per global item (k_item_ll)
    DevFlags* flags;
};

struct ThetaView {
    int n, resp0, n_local;    // local respondents [resp0, resp0 + n_local)
    int m_total, m_chunk, ld_chunk, world, C;
    const int8_t* y_rm;       // respondent-major rows of the local respondents: y_rm[il*ldm + J]
    long long ldm;
    const double* theta;      // n (current)
    double* theta_new;        // n
    const double* fstar_all;  // world chunks of NGRID x ld_chunk (grid-row-major: [idx*ld_chunk + jl])
    long long chunk_elems;
    const double* beta;       // 3 x m_total
    const double* thr;        // (C+1) x m_total item-major
    const double* table;      // optional [NGRID][m_total][C] table of cell terms (k_theta_table), or nullptr
    const float2* table32;    // optional [NGRID][m_total][C] table of (fp32 cell term, bound on its distance to the fp64 term) (k_theta_table32):
                              // proposals are screened on it, undecided ones are evaluated exactly (table, or on the fly when table == nullptr)
    int32_t* exact_evals;     // per respondent: evaluations that ran in fp64, or nullptr
    int y_in_smem;            // set by launch_ess_theta: the respondent's row of y fits the CTA's shared memory
    const double* theta_prior_sds;  // 2 x n
    int regime;               // 0: L_i = sqrt(1 + sd0_i^2) (CST/RDM), 1: sqrt(os^2 + 1e-6) (GP over time, H = 1)
    double theta_os;
    const double* z; long long z_stride;
    UniformSrc us;
    RngKey key;
    int32_t* evals;
    double* margin;
    DevFlags* flags;
};

// horizon > 1, GP-over-time regime (src/draw-theta.cpp:206-235): one H-dimensional ESS per respondent over the sessions' tables
struct ThetaSession {       // what the joint step reads / writes of one session (device pointers of that session's handle)
    const int8_t* y_rm;
    const double* theta;    // n (current)
    double* theta_new;      // n
    const double* fstar;    // NGRID x ld_chunk grid-row-major (world == 1)
    const double* beta;     // 3 x m
    const double* thr;      // (C+1) x m item-major
    const double* table;    // optional [NGRID][m][C] term table, or nullptr
};
constexpr int GPIRT_MAX_HORIZON = 32;
struct ThetaPanelView {
    int n, m, ld_chunk, C, H;
    long long ldm;
    const ThetaSession* ses;   // device array of H
    const double* L_time;      // H x H column-major lower factor (device)
    const double* z; long long z_stride;
    UniformSrc us;
    RngKey key;
    int32_t* evals;
    double* margin;
    DevFlags* flags;
};
cudaError_t launch_ess_theta_panel(cudaStream_t st, const ThetaPanelView& v);

// Item-sharded theta step (world > 1; draw_theta, src/draw-theta.cpp:26-84,112-170, with the per-respondent loop distributed over the ranks BY ITEMS):
// every rank keeps all n respondents and scores them against its own items only. The shrink sequence of an ESS unit under rejections is a function
// of its uniforms alone (the bracket shrinks to the rejected angle, no likelihood value enters), so a round evaluates the next THETA_BATCH
// proposals of every unfinished respondent at once: k_theta_partials writes this rank's partial log-likelihoods (slot 0: the current state,
// slots 1..THETA_BATCH: the proposals), the ranks exchange them (n x (THETA_BATCH + 1) doubles each), and k_theta_decide — the same arithmetic on
// every rank — adds the partials in rank order and takes the reference's decisions `ll' > log_y` proposal by proposal. No f* exchange, no
// replicated term table; the decisions are those of the single-GPU kernel unless a margin |ll' - log_y| is below fp64 summation noise (the
// recorded minimum margin tells).
constexpr int THETA_BATCH = 16;
struct ThetaItemsView {
    int n, m_local, C, world;
    const int8_t* y_rl;       // respondent-major rows over the LOCAL items: y_rl[i*ldml + j]
    long long ldml;
    const double* table;      // [NGRID][m_local][C] cell terms of the local items (k_theta_table)
    const double* theta;      // n (current)
    double* theta_new;        // n
    const double* theta_prior_sds;
    int regime; double theta_os;
    const double* z; long long z_stride;
    UniformSrc us;
    RngKey key;
    double* part_mine;        // this rank's partials [n][THETA_BATCH + 1]                                (k_theta_partials writes)
    const double* part_all;   // every rank's partials [world][n][THETA_BATCH + 1], chunk r = rank r      (k_theta_decide reads)
    int32_t* k_next;          // per respondent: proposals already decided (all rejected); -1 = finished
    double* ll0;              // per respondent: log-likelihood of the current state (sum over ranks, set in the first round)
    int32_t* n_active;        // device counter: respondents still unfinished after this round
    int32_t* evals; double* margin;
    DevFlags* flags;
};
cudaError_t launch_theta_partials(cudaStream_t st, const ThetaItemsView& v);
cudaError_t launch_theta_decide(cudaStream_t st, const ThetaItemsView& v);
// packs / unpacks this rank's chunk of (beta, thresholds, per-item ll) into one exchange buffer: [world][m_chunk * (3 + C + 1 + 1)] doubles
cudaError_t launch_pack_items(cudaStream_t st, const double* beta, const double* thr, const double* item_ll, int C1, int m_chunk, int rank, int world,
                              double* packed, int unpack);

enum ItemBlock { IB_F = 0, IB_BETA = 1, IB_THR = 2 };

cudaError_t launch_gen_normals(cudaStream_t st, double* out, int ld, int rows, int cols, RngKey key, uint32_t stream, uint32_t unit0);
cudaError_t launch_ess_items(cudaStream_t st, int block, const ItemView& v);
cudaError_t launch_item_ll(cudaStream_t st, const ItemView& v);
cudaError_t launch_ess_theta(cudaStream_t st, const ThetaView& v);
extern int g_theta_threads;
// table[(idx*m_total + J)*C + c-1] = log(Phi(tau_c - g) - Phi(tau_{c-1} - g) + 1e-6), g = f*(idx,J) + mu*(idx,J): every term the theta
// step can ask for, computed once per sweep instead of once per (respondent, proposal)
cudaError_t launch_theta_table(cudaStream_t st, const ThetaView& v, double* table);
// f[i + j*ldn] = fstar_chunk[grid_index(theta_i)*ld_chunk + j]  (src/gpirtMCMC.cpp:291-298)
cudaError_t launch_gather_f(cudaStream_t st, const double* theta, int n, const double* fstar_chunk, int ld_chunk, int m_local, double* f, int ldn);
// fstar_chunk[idx*ld_chunk + j] += mu_star(idx, item0 + j)  (src/draw-fstar.cpp:52)
cudaError_t launch_theta_table32(cudaStream_t st, const ThetaView& v, float2* table32);
cudaError_t launch_add_mustar(cudaStream_t st, double* fstar_chunk, int ld_chunk, int m_local, int item0, const double* beta, int p_beta);
cudaError_t launch_add_mustar_push(cudaStream_t st, double* fstar_chunk, int ld_chunk, int m_local, int item0, const double* beta, int p_beta,
                                   const PeerPush& peers);
cudaError_t launch_signal_peers(cudaStream_t st, const PeerPush& peers, unsigned long long epoch);
// only < 0: every peer's word must reach `epoch`; only >= 0: the word of that rank alone (a broadcast from one owner)
cudaError_t launch_wait_peers(cudaStream_t st, const unsigned long long* flags, int world, int rank, unsigned long long epoch, DevFlags* dev_flags, int only = -1);
// a rows x cols block (column-major, leading dimension ld, the same on every rank) copied to the same place of every peer's matrix: dst[p] points at
// the block's first element in peer p's memory (NVLink peer mapping). Rows start at a multiple of 2 of a 16-byte aligned matrix.
struct PeerBlockPush { int count; double* dst[MAX_PEERS]; };
cudaError_t launch_push_block(cudaStream_t st, const double* src, int ld, int rows, int cols, const PeerBlockPush& pb);
// sum of per-item values in index order (src/gpirtMCMC.cpp:325-331 accumulates item by item)
cudaError_t launch_ordered_sum(cudaStream_t st, const double* x, int n, double* out);
// layout converters between the reference's N x m column-major f* and the grid-row-major device table
cudaError_t launch_transpose(cudaStream_t st, const double* in, int ld_in, int rows, int cols, double* out, int ld_out);
cudaError_t launch_clamp_theta(cudaStream_t st, double* theta, int n);
cudaError_t launch_y_to_i8(cudaStream_t st, const double* y, long long count, int C, int8_t* out, DevFlags* flags);
cudaError_t launch_y_layouts(cudaStream_t st, const int8_t* y_full, int n, int m_total, int item0, int m_local, int8_t* y_im, int ldy,
                             int resp0, int n_local, int8_t* y_rm, long long ldm);

}  // namespace gpirt
