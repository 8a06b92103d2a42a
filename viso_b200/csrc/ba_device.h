// ba_device.h — device-side data layout and kernel launchers of the B200 bundle adjuster.
//
// Layout in HBM (one handle = one problem or one batch of independent problems):
//   observations  sorted by (point, pose): obs_cam int32[No], obs_uv float2[No]      12 B/obs
//                 (the owning point is implied by the CSR pt_ptr int32[Np+1])
//   points        double[2][Np][3]  (current / candidate, per-problem `cur` selects)
//   poses         double[2][Nc][7]  quaternion xyzw + t; cam_rt double[2][Nc][12] = R|t
//   per-obs rec   double[No][12] = {x/z, y/z, 1/z, w, G~(2x3), v~(2)}  written by linearize,
//                 gathered by schur_blocks and re-read by update (96 B, 32 B-sector aligned)
//   per-point     double[Np][9] = {L^-1 (6, lower), L^-1 b_l (3)}                     72 B
//   pair list     int2[Npairs] sorted by S block: the scatter-free gather order
//   S             packed lower-triangular 6x6 blocks per problem, double[nblocks][36]
#ifndef VISO_BA_DEVICE_H
#define VISO_BA_DEVICE_H

#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

namespace visoba {

// One-time per-DEVICE kernel configuration.  cudaFuncSetAttribute (dynamic shared memory above 48 KB, non-portable
// cluster sizes) applies to the current device only, and handles of one process may live on different devices and be
// driven from different host threads (viso_b200/stream.py), so the "configured" state is kept per device ordinal and
// behind a mutex instead of in a process-wide flag.
struct PerDeviceOnce {
    std::mutex mu;
    signed char state[64] = {}; // 0 unknown, 1 ok, -1 failed
    int n_sm[64] = {};
    template <typename F> bool run(F&& f)
    {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;
        std::lock_guard<std::mutex> lock(mu);
        if (state[dev] == 0) state[dev] = f() ? 1 : -1;
        return state[dev] == 1;
    }
    int sm_count()
    { // SMs of the current device
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 1;
        std::lock_guard<std::mutex> lock(mu);
        if (n_sm[dev] <= 0) {
            int n = 0;
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
            n_sm[dev] = n > 0 ? n : 1;
        }
        return n_sm[dev];
    }
};

constexpr int kRecDoubles = 12;   // per-observation record
constexpr int kPtInfoDoubles = 9; // per-point record
constexpr int kItemPart = 42;     // 36 block entries + 6 rhs entries per work item
constexpr int kChunkPoints = 1024;
constexpr int kItemMaxPairs = 512; // general path: most pairs in one work item of k_schur_blocks (one warp: 16 per lane)
constexpr int kSmallMaxN = 162; // dense per-problem solve in shared memory up to 27 free poses
// window fast path (ba_window.cu): one CTA per tile of consecutive points of one problem
constexpr int kWinThreads = 128;
// tile shape and residency of the linearise kernel: compile-time tuning knobs (tools/win_tune.sh sweeps them)
#ifndef VISO_WIN_OBS_CAP
#define VISO_WIN_OBS_CAP 640
#endif
#ifndef VISO_WIN_PT_CAP
#define VISO_WIN_PT_CAP 88
#endif
#ifndef VISO_WIN_LS_CTAS
#define VISO_WIN_LS_CTAS 3
#endif
constexpr int kWinObsCap = VISO_WIN_OBS_CAP; // observations per tile (their records live in shared memory)
constexpr int kWinPtCap = VISO_WIN_PT_CAP;   // points per tile
constexpr int kWinLsCtas = VISO_WIN_LS_CTAS; // CTAs of k_win_linearize_schur resident per SM (registers and shared memory)
constexpr int kWinMaxCams = 32; // keyframes per problem (visibility mask is 32 bits)
// per tile in win_blkptr: nblk + 1 offsets of the S blocks' pair lists, then nblk block indices in the ORDER the tile kernel
// takes them (longest list first: the four sub-warps of a warp then work on lists of similar length)
__host__ __device__ constexpr long long win_blkptr_entries(long long nblk) { return 2 * nblk + 1; }

struct WinTile { // one CTA's share of a problem: consecutive points and all their observations
    int32_t prob, p0, np, o0, no, cam0, ncam, nf;
    int64_t part_off;   // offset (doubles) of this tile's partial S blocks in tile_part
    int64_t blkptr_off; // first of the tile's nf (nf + 1) / 2 + 1 entries of win_blkptr
    int64_t list_off;   // first of the tile's entries of win_pairs
    int64_t pad;
};
static_assert(sizeof(WinTile) == 64, "WinTile is read as four int4");

struct LmState { // one per problem, lives in device memory; g2o OptimizationAlgorithmLevenberg state
    double lambda;
    double nu;
    double chi;         // robustified cost of the current state
    double chi_start;   // cost at the start of the current outer iteration
    double chi_initial; // cost before the first iteration
    double rho;
    double dp_sq;       // |delta_pose|^2 of the last solve (damping part of computeScale)
    int32_t cur;        // which of the two state buffers is current
    int32_t it;         // outer iterations finished
    int32_t q;          // trials spent in the current outer iteration
    int32_t trials;     // trials total
    int32_t active;     // 0 once finished or terminated
    int32_t terminated;
    int32_t ok2;        // last solve succeeded (S + lambda I was positive definite)
    int32_t accepted;   // a step was accepted in the current outer iteration
    int32_t target_it;  // outer iterations requested
    int32_t max_trials;
};

struct IterStat { // mirrors viso_ba_iter_stats
    double cost_before, cost_after, lambda;
    int32_t trials, accepted;
};

struct DevProblem {
    // sizes
    int32_t n_cams, n_points, n_problems;
    int32_t pt_begin, pt_end; // this rank's point range (sharding); whole range otherwise
    int64_t n_obs;
    int64_t n_pairs;
    int32_t n_items, n_diag_items, n_blocks, n_chunks;
    // camera model / robust kernel
    double fx, fy, cx, cy, huber_delta;
    // topology
    const int32_t* pt_ptr;
    const int32_t* pt_prob;
    const int32_t* obs_cam;
    const float2* obs_uv;
    const int32_t* cam_free;   // free index inside the problem or -1
    const int32_t* cam_prob;
    const int32_t* prob_cam0;  // [B+1]
    const int32_t* prob_tri0;  // [B+1] first S block of the problem
    const int32_t* prob_nfree; // [B]
    const int2* pairs;
    const int4* items;         // {pair_begin, pair_end, block, is_diag}
    const int32_t* diag_items; // indices of the items of diagonal blocks
    const int32_t* block_item_ptr; // [n_blocks+1]
    const int32_t* block_cam;  // [n_blocks] global cam of the block row if diagonal else -1
    const int32_t* diag_blk;   // [n_cams] S block of the pose's diagonal, -1 for fixed poses
    const int2* blk_ij;        // [n_blocks] (I, J) when only co-visible pairs are stored (large solver), else null:
                               // the blocks of a problem are then its dense lower triangle, block = tri0 + I (I + 1) / 2 + J
    const int4* chunks;        // {problem, pt_begin, pt_end, 0}
    const int32_t* prob_chunk_ptr; // [B+1]
    // window fast path: tiles, visibility masks, per-observation meta word
    int32_t n_tiles;
    const WinTile* tiles;
    const uint32_t* pt_mask;       // [Np] bit c: keyframe c of the point's problem observes it
    // observations once more, tile by tile and KEYFRAME-major inside a tile (slot order)
    const uint32_t* win_meta;      // [No] local keyframe | local point << 8
    const float2* win_uv;          // [No]
    const uint16_t* win_slot;      // [No] point-major observation -> slot inside its tile
    const uint32_t* win_blkptr;    // per tile: offsets of each S block's pair list (relative to list_off)
    const uint32_t* win_pairs;     // slot_a | slot_b << 16, per tile and S block, points in order
    const int32_t* prob_tile_ptr;  // [B+1]
    const int64_t* prob_part_off;  // [B+1] offset (doubles) of the problem's tile partials
    double* tile_part;             // per tile: [nblocks of its problem][kItemPart]
    double* tile_sums;             // [n_tiles][2]  {cost, scale} of the tile
    // straggler trials: CTA b works on tile tile_map[b] (the tiles of the problems still active), null = identity
    const int32_t* tile_map;
    int32_t n_map_tiles;
    // state
    double* pts[2];
    double* pose[2];
    double* cam_rt[2];
    LmState* lm;
    IterStat* iter_stats; // [target_it] (problem 0 only) or null
    // workspace
    double* rec;
    double* ptinfo;
    double* cost_pt;   // robust cost of the current state per point
    double* cost_new_pt;
    double* scale_pt;
    double* item_part;
    double* S_blk;     // [n_blocks][36] followed by bs
    double* bs;        // [n_cams][6]  (zero rows for fixed poses)
    double* dp;        // [n_cams][6]
    double* chunk_part;// [n_chunks][2]
    double* prob_sums; // [B][2]  (sharded: what gets all-reduced)
    unsigned long long* maxdiag_bits; // [B]
    int32_t* n_active; // [1]
    double lambda_scale; // 1 on the rank that adds lambda to diag(S) (rank 0), else 0
    // photometric edges (ba_photo.cu; reference bundle_adjuster.h:58-106); all null for a reprojection problem.  An edge's
    // obs_cam is its pose vertex / target keyframe, the source keyframe index rides in the bits of obs_uv.x
    const void* ph_frames;   // visoimg::DevImage [n_cams]: level-0 image of every pose's keyframe
    const double* ph_src_rt; // [n_cams][12] R|t the keyframes had at upload: srcFrame->Project never sees the optimiser's poses
    double* ph_sum;          // [n_obs][16] sums over the 16 samples of an edge (k_photo_sample)
    double* ph_mid;          // [n_obs][4]  w F C_tt F: the 2 x 2 middle of the edge's diagonal-block term
};

// ---- reduced camera system of ONE large problem (n > kSmallMaxN): LinearSolverDense, reference bundle_adjuster.h:113-114 ----
// The free poses are cut into P interior SEGMENTS separated by P - 1 SEPARATORS that are at least as wide as the
// co-visibility span (no S block couples two different interiors).  In the order [interiors ..., separators ...] the matrix
// is block-diagonal with a border; segment i is the partial Cholesky of
//     M_i = [ A_ii            ]   rows: interior i (padded to 64), then the LEFT separator's rows, then the RIGHT one's,
//           [ B_i   0 (fill)  ]   then the right-hand side as one more row
// over its first n_elim columns, all segments concurrently on disjoint CTA groups; what is left in the border block is the
// segment's contribution to the separator system T (block tridiagonal), which every CTA then factorises together; the
// interiors are back-substituted concurrently again.  P = 0: no partition, T is the whole matrix (dense / envelope).
constexpr int kCholBlock = 64; // every size below is a multiple of it (identity padding)
constexpr int kBarsPerTask = 4;  // counters per task of the large solve (ba_dense.cu: F, B, U, G)
constexpr int kCholDdStride = 64 * 64 + 128; // doubles per published diagonal factor: L_kk, 1 / diag, failure flag (ba_solve_plan.h: dd_doubles)

struct CholTask {
    double* A;              // row-major [nrows + 1][ld]; row nrows is the right-hand side; only the envelope is touched
    double* Dd;             // factorised diagonal blocks, [n_elim / 64][kCholDdStride] (A's own are stale)
    double* x;              // [nrows] solution; rows >= n_elim (the border) are inputs of the back substitution
    int32_t* flags;         // [nrows / 64] block of x published
    const int4* steps;      // [n_elim / 64] rows below the diagonal block that step k touches: [64 (k + 1), .x) and [.y, .z)
    const int32_t* first64; // [nrows / 64] first 64-column block of every 64-row block (what gets zeroed)
    int32_t ld, n_elim, nrows, n_real; // columns [n_real, n_elim) are identity padding
    int32_t cta0, ncta;     // the CTA group that factorises it
    int32_t sL, sR;         // border rows: [n_elim, n_elim + sL) left separator, [n_elim + sL, nrows) right separator
};

struct LargeSolve {
    int32_t n;              // 6 * free poses
    int32_t n_seg;          // interior segments P (0: unpartitioned)
    int32_t sep_cols;       // columns of one separator (multiple of 64)
    int32_t total_rows;     // sum over the tasks of nrows + 1 (grid of the zeroing kernel)
    const CholTask* tasks;  // device array [n_seg + 1]; the last one is T (the separator system, or the whole matrix)
    const int32_t* row_ptr; // [n_seg + 2] prefix of nrows + 1 over the tasks
    const int32_t* cam_task;// [free poses] task that eliminates the pose (separator poses: n_seg)
    const int32_t* cam_off; // [free poses] first column of the pose inside that task
    const int32_t* cam_sep; // [free poses] separator index of the pose or -1
    double* x;              // [n] the step, free-pose order
    int32_t* fail;          // [1]
    unsigned int* bars;     // [(n_seg + 1) * kBarsPerTask] counters of the tasks
};

// launchers (all asynchronous on `st`); each returns the number of kernels it launched
int launch_state_init(const DevProblem& P, int iterations, int max_trials, double lambda_init,
                      cudaStream_t st);
int launch_cam_rt(const DevProblem& P, int which, cudaStream_t st); // R|t of all poses of buffer `which` (< 0: each problem's current)
int launch_gather_current(const DevProblem& P, double* poses_out, double* pts_out, cudaStream_t st);
int launch_linearize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st);
int launch_schur_blocks(const DevProblem& P, int diag_only, cudaStream_t st);
int launch_block_finalize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st);
int launch_reduce(const DevProblem& P, int which, cudaStream_t st); // 0/2: cost_pt (2: even if inactive), 1: cost_new+scale
int launch_lm_init(const DevProblem& P, double tau, double lambda_init, cudaStream_t st);
int launch_solve_small(const DevProblem& P, int max_n, cudaStream_t st);
int launch_solve_large(const DevProblem& P, const LargeSolve& L, cudaStream_t st); // ba_dense.cu
int large_solve_max_n();                // largest n the dense solver takes
int large_solve_cta_count();            // CTAs of the cooperative solve kernel on the current device (one per SM)
int launch_update(const DevProblem& P, cudaStream_t st);
int launch_decide(const DevProblem& P, cudaStream_t st);
int launch_eval_cost(const DevProblem& P, cudaStream_t st); // cost of current state -> cost_pt
// photometric edges (ba_photo.cu)
int launch_photo_sample(const DevProblem& P, int which, int full, int even_if_inactive, cudaStream_t st);
int launch_photo_linearize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st);
int launch_photo_cost(const DevProblem& P, int which, int even_if_inactive, cudaStream_t st);
size_t photo_sum_doubles();
// window fast path
size_t win_shared_bytes();
int launch_win_linearize_schur(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st);
int launch_win_finalize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st);
int launch_win_update(const DevProblem& P, cudaStream_t st);
int launch_win_prob_sums(const DevProblem& P, int even_if_inactive, cudaStream_t st);
int launch_win_sums_decide(const DevProblem& P, cudaStream_t st); // prob_sums + decide of a trial in one launch
int launch_win_active_tiles(const DevProblem& P, int32_t* map, cudaStream_t st);

} // namespace visoba
#endif
