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
//        P  every CTA factorises the 64x64 diagonal block redundantly in shared memory (same
//           input, same order of operations -> same bits everywhere: no broadcast, no barrier),
//           then its warps solve their rows of the panel  X L_kk^T = A_panel  (one warp per row,
//           the row lives in registers, L_kk^T is read from shared memory without conflicts)
//        -- grid barrier --
//        U  64x64 tiles of the trailing lower triangle:  A_il -= X_i X_l^T
//        -- grid barrier --
//     CTA 0: backward substitution L^T x = y (the rhs rides along as row n of A, so the forward
//     substitution falls out of the factorisation).
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
constexpr size_t kCholSmemFixed = sizeof(double) * (3 * (size_t)kCB * kCS + 2 * kCB);

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

// one warp per S block: scatter the packed 6x6 blocks (and the rhs as row n) into the dense matrix
__global__ void __launch_bounds__(256) k_dense_assemble(DevProblem P, LargeSolve L)
{
    const int lane = threadIdx.x & 31;
    const int blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nf = P.prob_nfree[0];
    const int nblk = nf * (nf + 1) / 2;
    if (blk == 0 && lane == 0) *L.fail = 0;
    if (blk < nblk) {
        int I = (int)((sqrt(8.0 * (double)blk + 1.0) - 1.0) * 0.5);
        while ((I + 1) * (I + 2) / 2 <= blk) ++I;
        while (I * (I + 1) / 2 > blk) --I;
        const int J = blk - I * (I + 1) / 2;
        for (int e = lane; e < 36; e += 32) {
            const int r = e / 6, c = e - 6 * r;
            L.A[(size_t)(6 * I + r) * L.ld + 6 * J + c] = P.S_blk[36 * (size_t)blk + e];
        }
    } else if (blk < nblk + P.n_cams) {
        const int c = blk - nblk;
        const int f = P.cam_free[c];
        if (f >= 0 && lane < 6) L.A[(size_t)L.n * L.ld + 6 * f + lane] = P.bs[6 * (size_t)c + lane];
    }
}

__global__ void __launch_bounds__(kCT, 1) k_chol_coop(LargeSolve L)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) double smd[];
    double* const D = smd;                   // [kCB][kCS]  diagonal block while it is factorised
    double* const LT = D + kCB * kCS;        // [kCB][kCS]  LT[c][r] = L_kk[r][c]
    double* const T2 = LT + kCB * kCS;       // [kCB][kCS]  second tile of the trailing update
    double* const rinv = T2 + kCB * kCS;     // [kCB]       1 / L_kk[c][c]
    double* const xb = rinv + kCB;           // [kCB]
    double* const ysm = xb + kCB;            // [n]         CTA 0, backward substitution
    __shared__ int s_fail;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = L.n, ld = L.ld;
    const int G = gridDim.x;
    double* const A = L.A; // written and re-read across grid barriers: no __restrict__
    const int tx = tid & 15, ty = tid >> 4;
    if (tid == 0) s_fail = 0;

    for (int j0 = 0; j0 < n; j0 += kCB) {
        const int nb = min(kCB, n - j0);
        // ---------------- P: diagonal block, redundantly in every CTA ----------------
        for (int idx = tid; idx < kCB * kCB; idx += kCT) {
            const int r = idx >> 6, c = idx & 63;
            D[r * kCS + c] = (r < nb && c <= r) ? A[(size_t)(j0 + r) * ld + j0 + c] : 0.0;
        }
        __syncthreads();
        for (int j = 0; j < nb; ++j) {
            const double d = D[j * kCS + j];
            if (!(d > 0.0) || !isfinite(d)) { // every thread of every CTA sees the same d
                s_fail = 1;
                break;
            }
            const double l = sqrt(d);
            const double inv = 1.0 / l;
            // column j of L (into LT), rank-1 update of the rest of the lower triangle
            for (int i = j + 1 + tid; i < nb; i += kCT) LT[j * kCS + i] = D[i * kCS + j] * inv;
            if (tid == 0) LT[j * kCS + j] = l, rinv[j] = inv;
            for (int i = j + 1 + ty; i < nb; i += 16) {
                const double li = D[i * kCS + j] * inv;
                for (int k = j + 1 + tx; k <= i; k += 16) D[i * kCS + k] -= li * (D[k * kCS + j] * inv);
            }
            __syncthreads();
        }
        __syncthreads();
        if (s_fail) { // uniform over the whole grid: nobody reaches another barrier
            if (blockIdx.x == 0 && tid == 0) *L.fail = 1;
            return;
        }
        if (blockIdx.x == 0) { // the factor of the diagonal block, for the backward substitution
            double* dd = L.Dd + (size_t)(j0 / kCB) * kCB * kCB;
            for (int idx = tid; idx < kCB * kCB; idx += kCT) {
                const int r = idx >> 6, c = idx & 63;
                dd[idx] = (r < nb && c <= r) ? LT[c * kCS + r] : 0.0;
            }
        }
        // ---------------- P: panel rows, one warp per row (row n = rhs included) ----------------
        const int base = j0 + nb;
        for (int row = base + blockIdx.x * (kCT / 32) + warp; row <= n; row += G * (kCT / 32)) {
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
        grid.sync();
        // ---------------- U: trailing update over 64x64 tiles of the lower triangle ----------------
        const int rows_below = n + 1 - base;
        if (rows_below > 0) {
            const int nt = (rows_below + kCB - 1) / kCB;
            const int ntile = nt * (nt + 1) / 2;
            double* const Pi = D;   // the diagonal-block buffers are free again
            double* const Pl = T2;
            for (int t = blockIdx.x; t < ntile; t += G) {
                int ti = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
                while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
                while (ti * (ti + 1) / 2 > t) --ti;
                const int tl = t - ti * (ti + 1) / 2;
                const int i0 = base + kCB * ti, l0 = base + kCB * tl;
                __syncthreads();
                for (int idx = tid; idx < kCB * kCB; idx += kCT) {
                    const int r = idx >> 6, c = idx & 63;
                    Pi[r * kCS + c] = (i0 + r <= n && c < nb) ? A[(size_t)(i0 + r) * ld + j0 + c] : 0.0;
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
                        if (r <= n && c < n && c <= r) A[(size_t)r * ld + c] -= acc[i][j];
                    }
            }
        }
        grid.sync();
    }

    // ---------------- backward substitution  L^T x = y  (CTA 0; nobody syncs the grid any more) ----------------
    if (blockIdx.x != 0) return;
    for (int i = tid; i < n; i += kCT) ysm[i] = A[(size_t)n * ld + i];
    __syncthreads();
    for (int jb = ((n - 1) / kCB) * kCB; jb >= 0; jb -= kCB) {
        const int nb = min(kCB, n - jb);
        const double* dd = L.Dd + (size_t)(jb / kCB) * kCB * kCB;
        for (int idx = tid; idx < kCB * kCB; idx += kCT) D[(idx >> 6) * kCS + (idx & 63)] = dd[idx]; // D[r][c] = L[r][c]
        __syncthreads();
        if (warp == 0) {
            double y0 = lane < nb ? ysm[jb + lane] : 0.0;
            double y1 = lane + 32 < nb ? ysm[jb + lane + 32] : 0.0;
            for (int j = nb - 1; j >= 32; --j) {
                const double xj = shfl_d(y1, j - 32) / D[j * kCS + j];
                if (lane + 32 == j) y1 = xj;
                if (lane + 32 < j) y1 = fma(-xj, D[j * kCS + lane + 32], y1);
                y0 = fma(-xj, D[j * kCS + lane], y0);
            }
            for (int j = min(nb, 32) - 1; j >= 0; --j) {
                const double xj = shfl_d(y0, j) / D[j * kCS + j];
                if (lane == j) y0 = xj;
                if (lane < j) y0 = fma(-xj, D[j * kCS + lane], y0);
            }
            xb[lane] = y0, xb[lane + 32] = y1;
            if (lane < nb) ysm[jb + lane] = y0;
            if (lane + 32 < nb) ysm[jb + lane + 32] = y1;
        }
        __syncthreads();
        // y_above -= L[jb.., above]^T x_B : rows of L are contiguous in the column index
        for (int c = tid; c < jb; c += kCT) {
            double s0 = 0.0, s1 = 0.0;
            int j = 0;
            for (; j + 1 < nb; j += 2) {
                s0 = fma(A[(size_t)(jb + j) * ld + c], xb[j], s0);
                s1 = fma(A[(size_t)(jb + j + 1) * ld + c], xb[j + 1], s1);
            }
            if (j < nb) s0 = fma(A[(size_t)(jb + j) * ld + c], xb[j], s0);
            ysm[c] -= s0 + s1;
        }
        __syncthreads();
    }
    for (int i = tid; i < n; i += kCT) L.x[i] = ysm[i];
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

size_t large_solve_diag_doubles(int n) { return (size_t)((n + kCB - 1) / kCB) * kCB * kCB; }

int large_solve_max_n()
{ // the backward substitution keeps y in shared memory next to the three tiles
    return (int)((227 * 1024 - 1024 - kCholSmemFixed) / sizeof(double));
}

int launch_solve_large(const DevProblem& P, const LargeSolve& L, cudaStream_t st)
{
    static int n_sm = 0;
    static size_t configured = 0;
    const size_t smem = kCholSmemFixed + sizeof(double) * (size_t)L.n;
    if (n_sm == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
        if (n_sm <= 0) n_sm = 1;
    }
    if (smem > configured) {
        if (cudaFuncSetAttribute(k_chol_coop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
        configured = smem;
    }
    const int nf = L.n / 6;
    const long long warps = (long long)nf * (nf + 1) / 2 + P.n_cams;
    k_dense_assemble<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(P, L);
    // cooperative launch: the runtime guarantees that all CTAs are co-resident (grid barriers)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)n_sm), cfg.blockDim = dim3(kCT), cfg.dynamicSmemBytes = smem, cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, k_chol_coop, L);
    k_large_finish<<<1, 256, 0, st>>>(P, L);
    return 3;
}

} // namespace visoba
