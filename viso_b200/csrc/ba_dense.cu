// ba_dense.cu — K3 (large): dense solve of the reduced camera system of ONE problem whose
// n = 6 * (free poses) does not fit the shared-memory solver (n > kSmallMaxN), e.g. BASELINE
// config 3 (200 keyframes, n = 1194) and config 4 (2000 cameras, n = 11994).
//
// Reference: LinearSolverDense<PoseMatrixType> (reference include/bundle_adjuster.h:113-114):
// g2o copies the pose blocks into one dense matrix and runs a Cholesky-family solver on it.
//
// One PERSISTENT cooperative kernel (one CTA per SM, grid-wide barriers) replaces the former
// chain of 2 launches per 32-column panel:
//     for every 64-column block step
//        P  every CTA factorises the 64x64 diagonal block redundantly (same input, same order of
//           operations -> same bits everywhere: no broadcast, no barrier): 32x32 quarters, the
//           quarter on the diagonal by ONE warp in registers (shuffles, rsqrt on the pivot chain);
//           then its warps solve their rows of the panel  X L_kk^T = A_panel  (one warp per row,
//           the row lives in registers, L_kk^T is read from shared memory without conflicts)
//        -- grid barrier --
//        U  64x64 tiles of the trailing lower triangle:  A_il -= X_i X_l^T
//        -- grid barrier --
//     backward substitution L^T x = y (the rhs rides along as row n of A, so the forward
//     substitution falls out of the factorisation): block b of x is owned by one CTA, which folds
//     in every published x_b' (b' > b) and publishes x_b with a release flag; no grid barrier.
// Measured on B200 (tools/chol_bench.cu, n = 1194): 786 us, of which 240 us are the pivot chain
// (1194 columns x ~390 cycles of dependent FP64 latency) and 110 us the backward substitution.
// Everything is fp64; nothing is atomic; the summation order is fixed by the tiling, so two runs
// (and the ranks of a sharded solve, which all hold the same S) produce identical bits.
#include "ba_math.cuh"

#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace visoba {

namespace {

constexpr int kCB = 64;        // block size of the factorisation
constexpr int kCT = 256;       // threads per CTA
constexpr int kCS = kCB + 1;   // shared-memory row stride (doubles): odd -> column walks are conflict-free
constexpr size_t kCholSmem = sizeof(double) * (3 * (size_t)kCB * kCS + 7 * kCB);

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

#ifdef VISO_CHOL_PROFILE // tools/chol_bench.cu: time stamps of CTA 0 at the phase boundaries
__device__ unsigned long long g_chol_prof[16 * 512];
__device__ __forceinline__ void prof_mark(int step, int k)
{
    if (blockIdx.x == 0 && threadIdx.x == 0 && step < 512) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_chol_prof[16 * step + k] = t;
    }
}
#define PROF(step, k) prof_mark(step, k)
#else
#define PROF(step, k)
#endif

// Cholesky of a 32x32 block by one warp: lane i holds row i (lower part meaningful).  The serial
// chain of the whole factorisation runs through here (B200: DFMA 8.5, SHFL of a double 30, rsqrt 66
// cycles, tools/lat.cu).  Per column
//     l = r[j] * inv  ->  lane j+1 forms ITS next pivot locally (r[j+1] - l^2)  ->  one shuffle
//     broadcasts it  ->  rsqrt  (1/sqrt replaces sqrt + division),
// while column j of L reaches the other lanes through shared memory: one 16-byte broadcast load
// hands two entries to all 32 lanes, where shuffles cost two instructions per entry and made the
// warp shuffle-issue-bound.  `buf` = 2 x 32 doubles of shared memory owned by the calling warp.
// Returns true on a non-positive or non-finite pivot.
__device__ __forceinline__ bool warp_chol32(double (&r)[32], int lane, double* buf)
{
    double d = shfl_d(r[0], 0);
    bool bad = !(d > 0.0) || !(d < 1.7e308);
    double inv = rsqrt(d);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const double l = r[j] * inv; // lane j: L_jj = d / sqrt(d)
        r[j] = l;
        double* const col = buf + 32 * (j & 1);
        col[lane] = l;
        __syncwarp();
        double inv_next = 0.0;
        if (j + 1 < 32) {
            d = shfl_d(fma(-l, l, r[j + 1]), j + 1); // on lane j+1 this is exactly its updated diagonal
            bad |= !(d > 0.0) || !(d < 1.7e308);
            inv_next = rsqrt(d);
        }
#pragma unroll
        for (int k0 = (j + 1) & ~1; k0 < 32; k0 += 2) {
            const double2 lk = *reinterpret_cast<const double2*>(col + k0);
            if (k0 > j) r[k0] = fma(-l, lk.x, r[k0]);
            r[k0 + 1] = fma(-l, lk.y, r[k0 + 1]);
        }
        inv = inv_next;
    }
    return bad;
}

// zero the envelope of the dense matrix (row r: columns 64 * first64[r / 64] .. r); what lies left
// of the envelope is never read or written by the solver
__global__ void __launch_bounds__(256) k_dense_zero(LargeSolve L)
{
    const int r = blockIdx.x;
    const int c0 = kCB * L.first64[r / kCB];
    double* row = L.A + (size_t)r * L.ld;
    for (int c = c0 + threadIdx.x; c <= r; c += blockDim.x) row[c] = 0.0;
    if (r == 0 && threadIdx.x == 0) *L.fail = 0;
}

// one warp per stored S block: scatter the packed 6x6 blocks (and the rhs as row n) into the dense matrix
__global__ void __launch_bounds__(256) k_dense_assemble(DevProblem P, LargeSolve L)
{
    const int lane = threadIdx.x & 31;
    const int blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (blk < P.n_blocks) {
        const int2 ij = P.blk_ij[blk];
        for (int e = lane; e < 36; e += 32) {
            const int r = e / 6, c = e - 6 * r;
            L.A[(size_t)(6 * ij.x + r) * L.ld + 6 * ij.y + c] = P.S_blk[36 * (size_t)blk + e];
        }
    } else if (blk < P.n_blocks + P.n_cams) {
        const int c = blk - P.n_blocks;
        const int f = P.cam_free[c];
        if (f >= 0 && lane < 6) L.A[(size_t)L.n * L.ld + 6 * f + lane] = P.bs[6 * (size_t)c + lane];
    }
}

__global__ void __launch_bounds__(kCT, 1) k_chol_coop(LargeSolve L)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) double smd[];
    double* const D = smd;                   // [kCB][kCS]  tile buffer (trailing update, backward substitution)
    double* const LT = D + kCB * kCS;        // [kCB][kCS]  LT[c][r] = L_kk[r][c]
    double* const T2 = LT + kCB * kCS;       // [kCB][kCS]  second tile of the trailing update
    double* const rinv = T2 + kCB * kCS;     // [kCB]       1 / L_kk[c][c]
    double* const colb = rinv + kCB;         // [2][kCB]    column exchange of warp_chol32; y_b of the backward substitution
    double* const red = colb + 2 * kCB;      // [4][kCB]    partial sums of the backward substitution
    __shared__ int s_fail;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = L.n, ld = L.ld;
    const int G = gridDim.x;
    double* const A = L.A; // written and re-read across grid barriers: no __restrict__
    const int tx = tid & 15, ty = tid >> 4;
    const int nblk = (n + kCB - 1) / kCB;
    int* const flags = reinterpret_cast<int*>(L.Dd + (size_t)nblk * kCB * kCB); // [nblk] x_b published
    if (tid == 0) s_fail = 0;
    if (blockIdx.x == 0)
        for (int i = tid; i < nblk; i += kCT) flags[i] = 0; // visible to everybody after the first grid barrier

    for (int j0 = 0; j0 < n; j0 += kCB) {
        const int nb = min(kCB, n - j0);
        PROF(j0 / kCB, 0);
        // ---------------- P: diagonal block, redundantly in every CTA ----------------
        // 64x64 = [A11 .; A21 A22] in 32x32 quarters: warp 0 factorises A11 in registers (one row per
        // lane, shuffles only: no block barrier inside the column chain), all warps solve
        // L21 = A21 L11^-T and update A22 -= L21 L21^T, warp 0 factorises A22.
        for (int idx = tid; idx < kCB * kCB; idx += kCT) {
            const int rr = idx >> 6, c = idx & 63;
            D[rr * kCS + c] = (rr < nb && c <= rr) ? A[(size_t)(j0 + rr) * ld + j0 + c] : (rr == c ? 1.0 : 0.0);
        }
        __syncthreads();
        PROF(j0 / kCB, 1);
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            const int o = 32 * h;
            if (warp == 0) {
                double r[32];
#pragma unroll
                for (int k = 0; k < 32; ++k) r[k] = D[(o + lane) * kCS + o + k];
                const bool bad = warp_chol32(r, lane, colb);
                if (bad) s_fail = 1;
#pragma unroll
                for (int k = 0; k < 32; ++k) {
                    if (k <= lane) LT[(o + k) * kCS + o + lane] = r[k];
                    if (k == lane) rinv[o + k] = 1.0 / r[k];
                }
            }
            __syncthreads();
            if (h == 0) {
                PROF(j0 / kCB, 2);
                { // L21: rows 32..63, four per warp, the four chains interleaved
                    double a4[4];
#pragma unroll
                    for (int t = 0; t < 4; ++t) a4[t] = D[(32 + 4 * warp + t) * kCS + lane];
#pragma unroll 4
                    for (int c = 0; c < 32; ++c) {
                        const double ri = rinv[c], lc = LT[c * kCS + lane];
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const double xc = shfl_d(a4[t], c) * ri;
                            if (lane == c) a4[t] = xc;
                            if (lane > c) a4[t] = fma(-xc, lc, a4[t]);
                        }
                    }
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        D[(32 + 4 * warp + t) * kCS + lane] = a4[t];
                        LT[lane * kCS + 32 + 4 * warp + t] = a4[t];
                    }
                }
                __syncthreads();
                PROF(j0 / kCB, 3);
                { // A22 -= L21 L21^T, 2 x 2 entries per thread
                    double acc2[2][2] = {};
#pragma unroll 8
                    for (int k = 0; k < 32; ++k) {
                        const double u0 = D[(32 + ty) * kCS + k], u1 = D[(48 + ty) * kCS + k];
                        const double v0 = D[(32 + tx) * kCS + k], v1 = D[(48 + tx) * kCS + k];
                        acc2[0][0] = fma(u0, v0, acc2[0][0]), acc2[0][1] = fma(u0, v1, acc2[0][1]);
                        acc2[1][0] = fma(u1, v0, acc2[1][0]), acc2[1][1] = fma(u1, v1, acc2[1][1]);
                    }
#pragma unroll
                    for (int a = 0; a < 2; ++a)
#pragma unroll
                        for (int bb = 0; bb < 2; ++bb) {
                            const int rr = 32 + ty + 16 * a, c = 32 + tx + 16 * bb;
                            if (c <= rr) D[rr * kCS + c] -= acc2[a][bb];
                        }
                }
                __syncthreads();
                PROF(j0 / kCB, 4);
            }
        }
        if (s_fail) { // uniform over the whole grid: nobody reaches another barrier
            if (blockIdx.x == 0 && tid == 0) *L.fail = 1;
            return;
        }
        PROF(j0 / kCB, 5);
        if (blockIdx.x == 0) { // the factor of the diagonal block, for the backward substitution
            double* dd = L.Dd + (size_t)(j0 / kCB) * kCB * kCB;
            for (int idx = tid; idx < kCB * kCB; idx += kCT) {
                const int rr = idx >> 6, c = idx & 63;
                dd[idx] = (rr < nb && c <= rr) ? LT[c * kCS + rr] : 0.0;
            }
        }
        // ---------------- P: panel rows, one warp per row (row n = rhs included) ----------------
        // Only the rows inside the envelope take part: 64-row blocks up to last64[step]; below them the
        // block column is structurally zero and stays zero (no fill outside a monotone envelope).
        const int base = j0 + nb;
        const int rend = min(n, (L.last64[j0 / kCB] + 1) * kCB);
        const int nrow = max(rend - base, 0);
        for (int qr = blockIdx.x * (kCT / 32) + warp; qr <= nrow; qr += G * (kCT / 32)) {
            const int row = qr < nrow ? base + qr : n;
            double* a = A + (size_t)row * ld + j0;
            double a0 = lane < nb ? a[lane] : 0.0;
            double a1 = lane + 32 < nb ? a[lane + 32] : 0.0;
#pragma unroll 8
            for (int c = 0; c < 32; ++c) {
                if (c >= nb) break;
                const double xc = shfl_d(a0, c) * rinv[c];
                if (lane == c) a0 = xc;
                if (lane > c) a0 = fma(-xc, LT[c * kCS + lane], a0);
                a1 = fma(-xc, LT[c * kCS + lane + 32], a1);
            }
#pragma unroll 8
            for (int c = 32; c < kCB; ++c) {
                if (c >= nb) break;
                const double xc = shfl_d(a1, c - 32) * rinv[c];
                if (lane + 32 == c) a1 = xc;
                if (lane + 32 > c) a1 = fma(-xc, LT[c * kCS + lane + 32], a1);
            }
            if (lane < nb) a[lane] = a0;
            if (lane + 32 < nb) a[lane + 32] = a1;
        }
        PROF(j0 / kCB, 6);
        grid.sync();
        PROF(j0 / kCB, 7);
        // ---------------- U: trailing update over 64x64 tiles of the lower triangle ----------------
        { // tiles (ti >= tl) of the active rows, then one tile per tl for the rhs row n
            const int nt = (nrow + kCB - 1) / kCB;
            const int ntri = nt * (nt + 1) / 2;
            const int ntile = ntri + nt;
            double* const Pi = D;   // the diagonal-block buffers are free again
            double* const Pl = T2;
            for (int t = blockIdx.x; t < ntile; t += G) {
                const bool rhs = t >= ntri;
                int ti = 0, tl = t - ntri;
                if (!rhs) {
                    ti = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
                    while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
                    while (ti * (ti + 1) / 2 > t) --ti;
                    tl = t - ti * (ti + 1) / 2;
                }
                const int i0 = rhs ? n : base + kCB * ti, l0 = base + kCB * tl;
                const int ni = rhs ? 1 : min(kCB, n - i0); // valid rows of the tile
                __syncthreads();
                for (int idx = tid; idx < kCB * kCB; idx += kCT) {
                    const int r = idx >> 6, c = idx & 63;
                    Pi[r * kCS + c] = (r < ni && c < nb) ? A[(size_t)(i0 + r) * ld + j0 + c] : 0.0;
                    Pl[r * kCS + c] = (l0 + r < n && c < nb) ? A[(size_t)(l0 + r) * ld + j0 + c] : 0.0;
                }
                __syncthreads();
                double acc[4][4] = {};
#pragma unroll 4
                for (int k = 0; k < kCB; ++k) {
                    double av[4], bv[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) av[i] = Pi[(ty + 16 * i) * kCS + k];
#pragma unroll
                    for (int i = 0; i < 4; ++i) bv[i] = Pl[(tx + 16 * i) * kCS + k];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int r = i0 + ty + 16 * i, c = l0 + tx + 16 * j;
                        if (ty + 16 * i < ni && c < n && c <= r) A[(size_t)r * ld + c] -= acc[i][j];
                    }
            }
        }
        PROF(j0 / kCB, 8);
        grid.sync();
        PROF(j0 / kCB, 9);
    }

    // ---------------- backward substitution  L^T x = y, pipelined over the CTAs with flags ----------------
    // Block b of x belongs to CTA (nblk - 1 - b) mod G.  Its owner subtracts L[b', b]^T x_b' for every
    // b' > b as soon as x_b' is published (the L tile is already in registers by then), solves the
    // 64x64 triangle with one warp and publishes x_b.  Owners take their blocks top down and only
    // ever wait for higher blocks, and all CTAs are co-resident (cooperative launch): no deadlock.
    const int q = tid >> 6, cc = tid & 63;
    for (int b = nblk - 1 - (int)blockIdx.x; b >= 0; b -= G) {
        const int cb = b * kCB;
        const int nbb = min(kCB, n - cb);
        { // L_bb (row-major) and the reciprocals of its diagonal
            const double* dd = L.Dd + (size_t)b * kCB * kCB;
            for (int idx = tid; idx < kCB * kCB; idx += kCT) D[(idx >> 6) * kCS + (idx & 63)] = __ldcg(dd + idx);
        }
        double* const ysh = colb; // y_b (fetched before the waits)
        if (tid < kCB) ysh[tid] = tid < nbb ? __ldcg(A + (size_t)n * ld + cb + tid) : 0.0;
        double acc = 0.0;
        for (int bp = min(nblk - 1, L.last64[b]); bp > b; --bp) { // blocks below the envelope are zero
            const int rb = bp * kCB;
            const int nrow = min(kCB, n - rb);
            double t[16];
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
                const int j = q * 16 + jj;
                t[jj] = j < nrow ? __ldcg(A + (size_t)(rb + j) * ld + cb + cc) : 0.0;
            }
            if (tid == 0) {
                while (*reinterpret_cast<volatile int*>(flags + bp) == 0) {}
                __threadfence();
            }
            __syncthreads();
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
                const int j = q * 16 + jj;
                const double xj = j < nrow ? __ldcg(L.x + rb + j) : 0.0;
                acc = fma(t[jj], xj, acc);
            }
        }
        red[q * kCB + cc] = acc;
        __syncthreads();
        if (warp == 0) {
            // reciprocals of the diagonal, one pair per lane: the serial chain below multiplies
            const double ri0 = lane < nbb ? 1.0 / D[lane * kCS + lane] : 0.0;
            const double ri1 = lane + 32 < nbb ? 1.0 / D[(lane + 32) * kCS + lane + 32] : 0.0;
            double y0 = lane < nbb ? ysh[lane] - (red[lane] + red[kCB + lane] + red[2 * kCB + lane] + red[3 * kCB + lane]) : 0.0;
            double y1 = lane + 32 < nbb ? ysh[lane + 32] - (red[lane + 32] + red[kCB + lane + 32] + red[2 * kCB + lane + 32] + red[3 * kCB + lane + 32]) : 0.0;
            for (int j = nbb - 1; j >= 32; --j) {
                const double xj = shfl_d(y1, j - 32) * shfl_d(ri1, j - 32);
                if (lane + 32 == j) y1 = xj;
                if (lane + 32 < j) y1 = fma(-xj, D[j * kCS + lane + 32], y1);
                y0 = fma(-xj, D[j * kCS + lane], y0);
            }
            for (int j = min(nbb, 32) - 1; j >= 0; --j) {
                const double xj = shfl_d(y0, j) * shfl_d(ri0, j);
                if (lane == j) y0 = xj;
                if (lane < j) y0 = fma(-xj, D[j * kCS + lane], y0);
            }
            if (lane < nbb) L.x[cb + lane] = y0;
            if (lane + 32 < nbb) L.x[cb + lane + 32] = y1;
            __threadfence();
            __syncwarp();
            if (lane == 0) *reinterpret_cast<volatile int*>(flags + b) = 1;
        }
        __syncthreads();
    }
    PROF(nblk, 0);
}

__global__ void __launch_bounds__(256) k_large_finish(DevProblem P, LargeSolve L)
{
    __shared__ double sh[256];
    LmState& st = P.lm[0];
    if (!st.active) return;
    const int fail = *L.fail;
    const int cur = st.cur, cand = cur ^ 1;
    const int tid = threadIdx.x;
    for (int c = tid; c < P.n_cams; c += blockDim.x) {
        const int f = P.cam_free[c];
        double d[6];
        for (int r = 0; r < 6; ++r) d[r] = (f >= 0 && !fail) ? L.x[6 * f + r] : 0.0;
        for (int r = 0; r < 6; ++r) P.dp[6 * (size_t)c + r] = d[r];
        const double* pc = P.pose[cur] + 7 * (size_t)c;
        double* pn = P.pose[cand] + 7 * (size_t)c;
        if (f >= 0 && !fail) pose_oplus(d, pc, pn);
        else
            for (int r = 0; r < 7; ++r) pn[r] = pc[r];
        write_cam_rt(pn, P.cam_rt[cand] + 12 * (size_t)c);
    }
    double ss = 0.0;
    if (!fail)
        for (int i = tid; i < L.n; i += blockDim.x) ss += L.x[i] * L.x[i];
    sh[tid] = ss;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (tid < s) sh[tid] += sh[tid + s];
        __syncthreads();
    }
    if (tid == 0) {
        st.dp_sq = sh[0];
        st.ok2 = fail ? 0 : 1;
    }
}

} // namespace

size_t large_solve_diag_doubles(int n)
{ // the factorised diagonal blocks, followed by one int flag per block (backward substitution)
    const size_t nblk = (size_t)((n + kCB - 1) / kCB);
    return nblk * kCB * kCB + (nblk + 1) / 2;
}

int large_solve_max_n() { return 6 * 5000; } // 7.2 GB of dense lower triangle; nothing else limits n

// the factorisation + both substitutions alone (A assembled, x out); cooperative launch: the
// runtime guarantees that all CTAs are co-resident, which the grid barriers need
int launch_chol_coop(const LargeSolve& L, cudaStream_t st)
{
    static PerDeviceOnce once;
    const size_t smem = kCholSmem;
    if (!once.run([] { return cudaFuncSetAttribute(k_chol_coop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCholSmem) == cudaSuccess; })) return 0;
    const int n_sm = once.sm_count();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)n_sm), cfg.blockDim = dim3(kCT), cfg.dynamicSmemBytes = smem, cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, k_chol_coop, L) == cudaSuccess ? 1 : 0;
}

int launch_solve_large(const DevProblem& P, const LargeSolve& L, cudaStream_t st)
{
    const long long warps = (long long)P.n_blocks + P.n_cams;
    k_dense_zero<<<L.n, 256, 0, st>>>(L);
    k_dense_assemble<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(P, L);
    const int k = launch_chol_coop(L, st);
    k_large_finish<<<1, 256, 0, st>>>(P, L);
    return 3 + k;
}

} // namespace visoba
