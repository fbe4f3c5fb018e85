// Declarations of the dense fp64 building blocks (see linalg.cu).
#pragma once
#include <cmath>
#include <cuda_runtime.h>

#include <cstdint>
#include <vector>

#include "gemm.cuh"

namespace gpirt {

constexpr int CHOL_NB = 64;   // panel width of the blocked factorisation / solves
constexpr int LD_ALIGN = 16;  // leading dimensions are padded to this many doubles
inline int pad_ld(long long n) { return (int)((n + LD_ALIGN - 1) / LD_ALIGN * LD_ALIGN); }

cudaError_t build_K(cudaStream_t st, const double* x1, int n1, const double* x2, int n2, const double* sds_host3,
                    double* out, int ld, double diag_add);
cudaError_t build_K_time(cudaStream_t st, const double* x1, int n1, const double* x2, int n2, double os, double ls,
                         double t0, double t1, int kernel, double* out, int ld);

// In-place lower Cholesky of the n x n matrix A (lower triangle read, upper triangle zeroed on exit).
// *info_dev (device int, must be 0 on entry) receives the 1-based index of the first non-positive pivot.
// readers_ws: chol_ws_ints(n) device ints of workspace (zeroed by the call; one arrival counter per panel).
inline size_t chol_ws_ints(int n) { return (size_t)(n + CHOL_NB - 1) / CHOL_NB + 1; }
// Optional milestones: ev[c] is recorded on st as soon as rows [0, rows[c]) of L are final (right after the panel kernel that ends at or
// beyond rows[c]; rows ascending, the last one n), so a consumer of leading row blocks (nu = L Z) can start under the factorisation.
struct CholMilestones {
    int count = 0;
    int rows[6] = {0, 0, 0, 0, 0, 0};
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};
// Look-ahead for the two-level factorisation (n >= 4096): a helper stream and two events owned by the caller. With it the K = OB trailing update of
// an outer block is split by columns: the next outer block's columns are updated on `st`, whose latency-bound panel chain then runs while the rest
// of the update (the bulk of the flops) goes to the helper stream. Same tiles and k ranges as the single launch: bit-identical factor.
struct CholAux {
    cudaStream_t helper = nullptr;
    cudaEvent_t ev_panel = nullptr, ev_rest = nullptr;
};
// Persistent single-kernel factorisation (chol_persistent.cu): the task list of an n x n matrix, uploaded once per handle, and its state words.
struct CholPlan {
    const uint32_t* tasks = nullptr;  // device: chol_plan_tasks(n)
    int n_tasks = 0;
    int* state = nullptr;             // device: chol_plan_state_ints(n) ints
    int max_ctas = 0;
};
bool chol_is_persistent(int n, const CholPlan* plan);  // chol_lower_blocked(n, plan) runs the single persistent kernel: no row milestones before its end
int chol_plan_tiles(int n);
size_t chol_plan_state_ints(int n);
std::vector<uint32_t> chol_plan_tasks(int n);
// info_dev: 0, the failing leading minor (1-based), or -1 when the kernel gave up waiting (never on a healthy device)
cudaError_t chol_lower_persistent(cudaStream_t st, double* A, int ld, int n, int* info_dev, int* state_dev, const uint32_t* tasks_dev, int n_tasks,
                                  int max_ctas, long long* trace_dev = nullptr);
// Two-level factorisation (n >= 4096) distributed over `world` ranks that each hold the whole matrix: outer block b (256 columns) belongs to rank
// b % world. After the panels of an outer block every rank applies the K = 256 trailing update to ITS column blocks only (same tiles, same k order as
// the single launch: the factor is the one-GPU factor), the owner of the next outer block updates that block first and `push`es it (rows from its
// diagonal down) into every peer's matrix; the others `wait` for it before they factor its panels. Every rank ends with the whole L.
struct CholDist {
    int rank = 0, world = 1;
    void* ctx = nullptr;
    cudaError_t (*push)(void* ctx, cudaStream_t st, int block_id, double* block, int ld, int rows, int cols) = nullptr;
    cudaError_t (*wait)(void* ctx, cudaStream_t st, int block_id, int owner) = nullptr;
};
inline int chol_dist_blocks(int n) { return (n + 255) / 256; }  // exchange points a distributed factorisation of n columns may use (block ids 1 .. blocks-1)
// plan != nullptr and n below the two-level threshold: the persistent kernel; milestones are then all recorded at the end.
cudaError_t chol_lower_blocked(cudaStream_t st, double* A, int ld, int n, int* info_dev, int* readers_ws, const CholMilestones* marks = nullptr,
                               const CholAux* aux = nullptr, const CholPlan* plan = nullptr, const CholDist* dist = nullptr);
// B <- L^-1 B and B <- L^-T B for an n x nrhs right-hand side, in place.
cudaError_t trsm_left_lower(cudaStream_t st, const double* L, int ldl, int n, double* B, int ldb, int nrhs);
cudaError_t trsm_left_lower_t(cudaStream_t st, const double* L, int ldl, int n, double* B, int ldb, int nrhs);

// Linv <- L^-1 (lower triangular, strict upper triangle of Linv must already be zero); scratch holds trtri_scratch_elems(n) doubles.
size_t trtri_scratch_elems(int n);
// dist: the merge levels with blocks of 512 columns or more (all but 1/8 of the flops at config-4 sizes) are split by COLUMNS of every pair's block over
// the ranks — both products of a level are column-wise independent — at 64-column tile boundaries placed so that the two products' work together is even
// (the first one is triangular in its columns); every rank `push`es its columns of the new blocks into its peers' Linv and all `sync` once per level.
struct TrtriDist {
    int rank = 0, world = 1;
    void* ctx = nullptr;
    cudaError_t (*push)(void* ctx, cudaStream_t st, double* block, int ld, int rows, int cols) = nullptr;
    cudaError_t (*sync)(void* ctx, cudaStream_t st) = nullptr;
};
cudaError_t trtri_lower(cudaStream_t st, const double* L, int ld, int n, double* Linv, int ldi, double* scratch, const TrtriDist* dist = nullptr);

// ---- who computes what in the distributed theta-only chain: pure functions of (sizes, world, rank), the same on every rank (tests/test_dist_plan.py) ----
inline int chol_dist_owner(int block, int world) { return block % world; }  // 256-column outer block `block` of the two-level Cholesky
inline bool trtri_dist_level(int b, int world) { return world > 1 && b >= 512 && b / 64 >= world; }  // is the merge level with b-column blocks split?
// first column of rank r's share of a b-column block of a split level (r = world: b). Work per column at relative position x: 2(1 - x) in the first
// product (A^-1 is lower triangular) + 1 in the second; the cumulative 3x - x^2 is cut into `world` equal parts, at whole 64-column tiles.
inline int trtri_dist_cut(int b, int world, int r) {
    if (r <= 0) return 0;
    if (r >= world) return b;
    const double x = (3.0 - sqrt(9.0 - 8.0 * (double)r / world)) * 0.5;
    const int c = (int)(x * (b / 64) + 0.5) * 64;
    return c < 0 ? 0 : (c > b ? b : c);
}
// columns [lo, hi) of the N grid points whose L^-1 k* and K^-1 k* rank r forms: runs of whole 64-column tiles
inline void kstar_dist_cols(int N, int world, int r, int* lo, int* hi) {
    const int tiles = (N + 63) / 64, per = (tiles + world - 1) / world;
    *lo = r * per * 64 < N ? r * per * 64 : N;
    *hi = (r + 1) * per * 64 < N ? (r + 1) * per * 64 : N;
}

}  // namespace gpirt
