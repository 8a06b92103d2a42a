// ba_dense.cu — K3 (large): solve of the reduced camera system of ONE problem whose n = 6 * (free poses) does not fit
// the shared-memory solver (n > kSmallMaxN), e.g. BASELINE config 3 (200 keyframes, n = 1194, dense) and config 4
// (2000 cameras, n = 11994, banded: co-visibility spans ~25 cameras).
//
// Reference: LinearSolverDense<PoseMatrixType> (reference include/bundle_adjuster.h:113-114): g2o copies the pose blocks
// into one dense matrix and runs a Cholesky-family solver on it, on one core.
//
// Here: ONE persistent cooperative kernel (one CTA per SM) that works through a list of CholTask (ba_device.h):
//   phase 1  every interior SEGMENT of the pose chain is eliminated by its own group of CTAs, all segments at once
//            (group barriers: a counter in global memory, the launch is cooperative so every CTA is resident);
//   phase 2  what the segments leave in their border blocks is added, in a fixed order, to the separator system T;
//   phase 3  T (block tridiagonal; or, unpartitioned, the whole matrix) is factorised and solved by all CTAs;
//   phase 4  the interiors are back-substituted, all segments at once again.
// A banded system of n = 11994 thus costs ~21 + 21 block steps instead of 188 sequential ones, and the same plan is what
// every rank of a sharded solve executes redundantly on the all-reduced S (identical bits on every rank, no broadcast).
//
// One 64-column block step of a task (the unit everything is padded to):
//     D  every CTA of the group factorises the 64x64 diagonal block redundantly in shared memory (same input, same
//        order of operations -> same bits: no broadcast, no barrier): 8x8 blocks, the diagonal one by ONE thread in
//        registers (the serial chain of the whole factorisation: rsqrt + two dependent FMAs per pivot, no shuffles, no
//        barrier inside), its panel by one thread per row, the 56x56 remainder by all 256 threads;
//     P  panel  X L_kk^T = A_panel: one warp per FOUR rows (the four chains interleave), the row in registers;
//        -- group barrier --
//     U  64x64 tiles of the trailing lower triangle over the step's row ranges: A_il -= X_i X_l^T;
//        -- group barrier --
// The right-hand side rides along as one more row of every task, so the forward substitution falls out of the
// factorisation.  Backward substitution L^T x = y: block b of x is owned by one CTA of the group, which folds in every
// published x_b' (b' > b) and publishes x_b with a release flag; no barrier.
// Everything is fp64; nothing is atomic on floating-point data; the summation order is fixed by the tiling.
#include "ba_math.cuh"

#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace visoba {

namespace {

constexpr int kCB = kCholBlock; // 64: block size of the factorisation
constexpr int kCT = 256;        // threads per CTA
constexpr int kPR = 1;         // panel rows per warp (their substitution chains interleave)
constexpr int kCS = kCB + 1;    // shared-memory row stride (doubles): odd -> column walks are conflict-free
constexpr size_t kCholSmem = sizeof(double) * (3 * (size_t)kCB * kCS + 7 * kCB);

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

#ifdef VISO_CHOL_PROFILE // tools/chol_bench.cu: time stamps of CTA 0 at the phase boundaries of the LAST task (T)
__device__ unsigned long long g_chol_prof[16 * 512];
__device__ __forceinline__ void prof_mark(bool on, int step, int k)
{
    if (on && threadIdx.x == 0 && step < 512) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_chol_prof[16 * step + k] = t;
    }
}
#define PROF(on, step, k) prof_mark(on, step, k)
__device__ long long g_c64_prof[4 * 8]; // cycles of the sub-step phases of the last cta_chol64 of CTA 0
#define C64(i, v) ((blockIdx.x == 0 && threadIdx.x == 0) ? (void)(g_c64_prof[i] = (v)) : (void)0)
#else
#define PROF(on, step, k)
#define C64(i, v) ((void)0)
#endif

// ---- barrier of a CTA group: monotone counter, target = visits * group size (zeroed before the launch) ----
__device__ __forceinline__ void group_barrier(unsigned int* ctr, unsigned int target)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        unsigned int v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        } while (v < target);
        __threadfence();
    }
    __syncthreads();
}

// 8x8 diagonal block at (c0, c0) of D, by ONE thread in registers: the pivot chain of the whole factorisation
__device__ __forceinline__ void chol8_thread(double* D, double* rinv, int c0, int* s_fail)
{
    double a[36];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) a[i * (i + 1) / 2 + j] = D[(c0 + i) * kCS + c0 + j];
    bool bad = false;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const double d = a[j * (j + 1) / 2 + j];
        bad |= !(d > 0.0) || !(d < 1.7e308);
        const double inv = rsqrt(d);
        a[j * (j + 1) / 2 + j] = d * inv;
        rinv[c0 + j] = inv;
#pragma unroll
        for (int i = j + 1; i < 8; ++i) a[i * (i + 1) / 2 + j] *= inv;
#pragma unroll
        for (int i = j + 1; i < 8; ++i)
#pragma unroll
            for (int k = j + 1; k <= i; ++k) a[i * (i + 1) / 2 + k] = fma(-a[i * (i + 1) / 2 + j], a[k * (k + 1) / 2 + j], a[i * (i + 1) / 2 + k]);
    }
    if (bad) *s_fail = 1;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) D[(c0 + i) * kCS + c0 + j] = a[i * (i + 1) / 2 + j];
}

// named barriers of the CTA (id 0 is __syncthreads): a warp that has nothing to wait for ARRIVES and goes on
__device__ __forceinline__ void bar_arrive(int id, int count)
{
    __threadfence_block();
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// Cholesky of the 64x64 block in D (lower triangle, row stride kCS; the upper triangle is and stays zero) by the whole
// CTA, in place; rinv[c] = 1 / L[c][c].  Returns true on a non-positive or non-finite pivot (uniform over the CTA).
// Eight sub-steps of 8 columns.  Warp 0 carries the pivot chain: its lanes 0..7 solve the panel rows of the NEXT diagonal
// block, it arrives at the panel barrier without waiting, updates that block (36 entries) and its lane 0 factorises it
// in registers, while warps 1..7 wait for the whole panel and apply  remainder -= panel panel^T  in 2 x 2 micro-tiles.
// The single-thread pivot chain thus runs beside the parallel work, and the chain never waits for a full-CTA barrier
// between a panel and the diagonal block that depends on it.
__device__ __forceinline__ bool cta_chol64(double* D, double* rinv, int* s_fail, int tid)
{
#ifdef VISO_CHOL_PROFILE
    long long pq0 = clock64(), pq1, pq2, pq3;
#endif
    if (tid == 0) chol8_thread(D, rinv, 0, s_fail);
    __syncthreads();
    C64(28, clock64() - pq0);
#pragma unroll 1
    for (int c0 = 0; c0 < kCB - 8; c0 += 8) {
        const int m = kCB - c0 - 8; // rows below the 8x8 block
#ifdef VISO_CHOL_PROFILE
        pq0 = clock64();
#endif
        if (tid < m) { // panel: one thread per row, x L88^T = a
            const int r = c0 + 8 + tid;
            double x[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = D[r * kCS + c0 + j];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                double s = x[j];
#pragma unroll
                for (int t = 0; t < j; ++t) s = fma(-x[t], D[(c0 + j) * kCS + c0 + t], s);
                x[j] = s * rinv[c0 + j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) D[r * kCS + c0 + j] = x[j];
        }
#ifdef VISO_CHOL_PROFILE
        pq1 = clock64();
#endif
        if (tid < 32) { // the next diagonal block (36 entries: lanes 0..3 take two), then its factorisation by lane 0
            __syncwarp(); // its panel rows are rows 0..7 of this warp
            bar_arrive(1, kCT);
#ifdef VISO_CHOL_PROFILE
            const long long pa = clock64();
#endif
            // entry e = tid of the packed lower triangle: row rr, column cc (no loop: this is the pivot chain's time)
            const int rr = tid >= 28 ? 7 : tid >= 21 ? 6 : tid >= 15 ? 5 : tid >= 10 ? 4 : tid >= 6 ? 3 : tid >= 3 ? 2 : tid >= 1 ? 1 : 0;
            const int cc = tid - rr * (rr + 1) / 2; // tid = 28..31 are the first entries of row 7; entries 32..35 the last four
            const bool two = tid < 4;
            const double* pr = D + (c0 + 8 + rr) * kCS + c0;
            const double* pc = D + (c0 + 8 + cc) * kCS + c0;
            const double* pr2 = D + (c0 + 8 + 7) * kCS + c0;
            const double* pc2 = D + (c0 + 8 + 4 + (tid & 3)) * kCS + c0;
            // two half-sums per entry: the dependent FMA chain is 4 + 1 long instead of 8
            double sa = 0.0, sb = 0.0, s2a = 0.0, s2b = 0.0;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                sa = fma(pr[t], pc[t], sa), sb = fma(pr[t + 4], pc[t + 4], sb);
                s2a = fma(pr2[t], pc2[t], s2a), s2b = fma(pr2[t + 4], pc2[t + 4], s2b);
            }
            D[(c0 + 8 + rr) * kCS + c0 + 8 + cc] -= sa + sb;
            if (two) D[(c0 + 8 + 7) * kCS + c0 + 8 + 4 + tid] -= s2a + s2b;
            __syncwarp();
#ifdef VISO_CHOL_PROFILE
            const long long pb = clock64();
            if (c0 == 48) C64(30, pa - pq1), C64(31, pb - pa);
#endif
            if (tid == 0) chol8_thread(D, rinv, c0 + 8, s_fail);
        } else {
            bar_sync(1, kCT); // every panel row is written
            // remainder -= panel panel^T below the next diagonal block, in 2x2 micro-tiles of the lower triangle: tile rows
            // R = 4 .. m/2 - 1 hold R + 1 tiles each; row 4 + rp is folded with row m/2 - 1 - rp into one line of
            // W = m/2 + 5 tiles, so the enumeration is rectangular and no lane idles on the upper triangle
            const int h = m >> 1, W = h + 5, nline = (h - 4) >> 1;
            for (int idx = tid - 32; idx < nline * W; idx += kCT - 32) {
                const int rp = idx / W, q = idx - rp * W;
                int R = 4 + rp, C = q;
                if (q > R) R = h - 1 - rp, C = q - (5 + rp);
                const double* pr0 = D + (c0 + 8 + 2 * R) * kCS + c0;
                const double* pc0 = D + (c0 + 8 + 2 * C) * kCS + c0;
                double s00 = 0.0, s01 = 0.0, s10 = 0.0, s11 = 0.0;
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const double r0 = pr0[t], r1 = pr0[kCS + t], q0 = pc0[t], q1 = pc0[kCS + t];
                    s00 = fma(r0, q0, s00), s01 = fma(r0, q1, s01), s10 = fma(r1, q0, s10), s11 = fma(r1, q1, s11);
                }
                double* o = D + (c0 + 8 + 2 * R) * kCS + c0 + 8 + 2 * C;
                o[0] -= s00;
                if (C < R) o[1] -= s01;
                o[kCS] -= s10, o[kCS + 1] -= s11;
            }
        }
#ifdef VISO_CHOL_PROFILE
        pq2 = clock64();
#endif
        __syncthreads();
#ifdef VISO_CHOL_PROFILE
        pq3 = clock64();
        C64(c0 / 8 * 4 + 0, pq1 - pq0), C64(c0 / 8 * 4 + 1, pq2 - pq1), C64(c0 / 8 * 4 + 2, pq3 - pq2);
#endif
    }
    return *s_fail != 0;
}

// The factor of a diagonal block in T.Dd: L_kk row-major [64][64] (straight from the factorised block: its upper triangle
// is zero), then rinv[64], then the failure flag of the factorisation so far (0 / 1), padded to kCholDdStride doubles.
__device__ __forceinline__ void publish_factor(double* dd, const double* D, const double* rinv, bool bad, int tid)
{
    for (int idx = tid; idx < kCB * kCB; idx += kCT) dd[idx] = D[(idx >> 6) * kCS + (idx & 63)];
    if (tid < kCB) dd[kCB * kCB + tid] = rinv[tid];
    if (tid == kCB) dd[kCB * kCB + kCB] = bad ? 1.0 : 0.0;
}

__device__ __forceinline__ bool load_factor(const double* dd, double* LT, double* rinv, int tid)
{
    for (int idx = tid; idx < kCB * kCB; idx += kCT) LT[(idx & 63) * kCS + (idx >> 6)] = __ldcg(dd + idx);
    if (tid < kCB) rinv[tid] = __ldcg(dd + kCB * kCB + tid);
    const bool bad = __ldcg(dd + kCB * kCB + kCB) != 0.0;
    __syncthreads();
    return bad;
}

// row q of a step's row list: the band range, then the border range, then the right-hand side
__device__ __forceinline__ int step_row(int q, int base, int n1, int r2b, int n2, int rhs)
{
    return q < n1 ? base + q : (q < n1 + n2 ? r2b + (q - n1) : rhs);
}

// ---- counters of a task (L.bars + kBarsPerTask * task, zeroed before the launch; all monotone) ----
//   [0] F  factors published by the lead CTA of the set           (k + 1 after the factor of diagonal block k)
//   [1] B  worker arrivals after the panel of a step              ((k + 1) * workers after step k)
//   [2] U  worker arrivals after the trailing update of a step    ((k + 1) * workers after step k)
//   [3] G  plain barrier of the set (before the back substitution of an interior segment)
__device__ __forceinline__ void ctr_arrive(unsigned int* ctr)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
    }
}

__device__ __forceinline__ void ctr_wait(const unsigned int* ctr, unsigned int target)
{
    if (threadIdx.x == 0) {
        unsigned int v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        } while (v < target);
        __threadfence();
    }
    __syncthreads();
}

// panel rows  X L_kk^T = A_panel  of one step: one warp per R rows (their substitution chains interleave), the row in
// registers, lane = column (and column + 32).  x of a finished column lives in xv (off the dependent chain: shuffle ->
// multiply -> FMA -> next shuffle); lanes left of the column see L^T entries that are zero and keep their values.
template <int R>
__device__ __forceinline__ void panel_rows(double* A, int ld, int j0, int base, int n1, int r2b, int n2, int nrows, int qb, int nrow,
                                           int first_warp, int n_warps, const double* LT, const double* rinv, int lane)
{ // rows qb .. nrow - 1 of the step's row list
    for (int q0 = qb + R * first_warp; q0 < nrow; q0 += R * n_warps) {
        double* a[R];
        double a0[R], a1[R], xv0[R], xv1[R];
#pragma unroll
        for (int t = 0; t < R; ++t) {
            const int q = min(q0 + t, nrow - 1); // a short last group repeats its last row (not written back)
            a[t] = A + (size_t)step_row(q, base, n1, r2b, n2, nrows) * ld + j0;
            a0[t] = __ldcg(a[t] + lane), a1[t] = __ldcg(a[t] + lane + 32);
            xv0[t] = 0.0, xv1[t] = 0.0;
        }
#pragma unroll 4
        for (int c = 0; c < 32; ++c) {
            const double ri = rinv[c], l0 = LT[c * kCS + lane], l1 = LT[c * kCS + lane + 32];
            double xs[R];
#pragma unroll
            for (int t = 0; t < R; ++t) xs[t] = shfl_d(a0[t], c);
#pragma unroll
            for (int t = 0; t < R; ++t) {
                const double xc = xs[t] * ri;
                if (lane == c) xv0[t] = xc;
                a0[t] = fma(-xc, l0, a0[t]);
                a1[t] = fma(-xc, l1, a1[t]);
            }
        }
#pragma unroll 4
        for (int c = 32; c < kCB; ++c) {
            const double ri = rinv[c], l1 = LT[c * kCS + lane + 32];
            double xs[R];
#pragma unroll
            for (int t = 0; t < R; ++t) xs[t] = shfl_d(a1[t], c - 32);
#pragma unroll
            for (int t = 0; t < R; ++t) {
                const double xc = xs[t] * ri;
                if (lane + 32 == c) xv1[t] = xc;
                a1[t] = fma(-xc, l1, a1[t]);
            }
        }
#pragma unroll
        for (int t = 0; t < R; ++t)
            if (q0 + t < nrow) a[t][lane] = xv0[t], a[t][lane + 32] = xv1[t];
    }
}

// Partial Cholesky of a task: eliminate its first n_elim columns.  gb / G: this CTA's index in, and the size of, the
// CTA set that works on the task; bars: the task's counters.  Returns false when a pivot failed (uniform over the set).
//
// Roles.  CTA 0 of the set is the LEAD: it owns the pivot chain.  For every step k it waits until the panel of step k is
// written (counter B), forms the next diagonal block  A_{k+1,k+1} - X_{k+1,k} X_{k+1,k}^T  in shared memory, factorises
// it and publishes the factor (counter F).  The other CTAs are WORKERS: wait for factor k, panel of step k, (B), trailing
// update of step k without the lead's tile, (U).  The lead's 64-pivot chain thus runs beside the workers' update instead
// of in front of it, and no CTA repeats it.  A set of one CTA plays both roles in turn.
// (Tried and dropped: the lead solving the next row block itself in shared memory — 8.5 us per step for the 64-row solve
// against 5.3 us of waiting for the workers' panel: C3 491 -> 530 us.)
__device__ bool task_factor(const CholTask& T, int gb, int G, unsigned int* bars, double* smd, int* s_fail, bool prof)
{
    double* const D = smd;                // [kCB][kCS] diagonal block, then tile buffer of the trailing update
    double* const LT = D + kCB * kCS;     // [kCB][kCS] LT[c][r] = L_kk[r][c]
    double* const T2 = LT + kCB * kCS;    // [kCB][kCS] second tile of the trailing update
    double* const rinv = T2 + kCB * kCS;  // [kCB]      1 / L_kk[c][c]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ld = T.ld, nrows = T.nrows;
    double* const A = T.A; // written and re-read across the counters by other CTAs: L2 loads (__ldcg), no __restrict__
    const int tx = tid & 15, ty = tid >> 4;
    const int nsteps = T.n_elim / kCB;
    const bool lead = gb == 0, worker = gb > 0 || G == 1;
    const int nW = G > 1 ? G - 1 : 1, w = G > 1 ? gb - 1 : 0;
    unsigned int* const cF = bars, * const cB = bars + 1, * const cU = bars + 2;
    bool ok = true;
    if (lead) { // the factor of diagonal block 0
        for (int idx = tid; idx < kCB * kCB; idx += kCT) {
            const int rr = idx >> 6, c = idx & 63;
            D[rr * kCS + c] = c <= rr ? __ldcg(A + (size_t)rr * ld + c) : 0.0;
        }
        __syncthreads();
        if (cta_chol64(D, rinv, s_fail, tid)) ok = false;
        publish_factor(T.Dd, D, rinv, !ok, tid);
        ctr_arrive(cF);
    }
    for (int k = 0; k < nsteps; ++k) {
        const int j0 = k * kCB;
        const int4 st = __ldg(T.steps + k);
        const int base = j0 + kCB;
        const int n1 = max(st.x - base, 0), r2b = st.y, n2 = max(st.z - st.y, 0);
        const int n1t = n1 / kCB, nt = n1t + n2 / kCB;
        const int ntri = nt * (nt + 1) / 2, ntile = ntri + nt;
        const bool next = k + 1 < nsteps;
        const bool lead_tile = next && n1t >= 1; // tile 0 of the update is the next diagonal block: the lead's
        const unsigned int done = (unsigned)(k + 1) * (unsigned)nW;
        if (worker) {
            PROF(prof && gb == 1, k, 8);
            // ---------------- P: panel of step k ----------------
            ctr_wait(cF, (unsigned)k + 1u);
            if (load_factor(T.Dd + (size_t)k * kCholDdStride, LT, rinv, tid)) ok = false;
            PROF(prof && gb == 1, k, 9);
            const int nrow = n1 + n2 + 1, nwarps = nW * (kCT / 32), fw = w * (kCT / 32) + warp;
            const int qb = 0;
            if (nrow - qb > 2 * nwarps) panel_rows<4>(A, ld, j0, base, n1, r2b, n2, nrows, qb, nrow, fw, nwarps, LT, rinv, lane);
            else if (nrow - qb > nwarps) panel_rows<2>(A, ld, j0, base, n1, r2b, n2, nrows, qb, nrow, fw, nwarps, LT, rinv, lane);
            else panel_rows<1>(A, ld, j0, base, n1, r2b, n2, nrows, qb, nrow, fw, nwarps, LT, rinv, lane);
            PROF(prof && gb == 1, k, 10);
            ctr_arrive(cB);
        }
        if (lead && next) {
            // ---------------- D: the next diagonal block, its factor ----------------
            PROF(prof, k, 0);
            double dg[4][4]; // the block's own entries: final once the update of step k - 1 is through
            if (k > 0) ctr_wait(cU, done - (unsigned)nW);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = ty + 16 * i, c = tx + 16 * j;
                    dg[i][j] = c <= r ? __ldcg(A + (size_t)(base + r) * ld + base + c) : 0.0;
                }
            PROF(prof, k, 1);
            double acc[4][4] = {};
            if (lead_tile) {
                ctr_wait(cB, done);
                PROF(prof, k, 2);
                double* const Pi = D;
                double v[16];
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const int idx = tid + q * kCT;
                    v[q] = __ldcg(A + (size_t)(base + (idx >> 6)) * ld + j0 + (idx & 63));
                }
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const int idx = tid + q * kCT;
                    Pi[(idx >> 6) * kCS + (idx & 63)] = v[q];
                }
                __syncthreads();
#pragma unroll 4
                for (int kk = 0; kk < kCB; ++kk) { // lower micro-tiles only (j <= i): the block is symmetric
                    double av[4], bv[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) av[i] = Pi[(ty + 16 * i) * kCS + kk];
#pragma unroll
                    for (int i = 0; i < 4; ++i) bv[i] = Pi[(tx + 16 * i) * kCS + kk];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j <= i; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
                }
                __syncthreads();
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = ty + 16 * i, c = tx + 16 * j;
                    D[r * kCS + c] = c <= r ? dg[i][j] - acc[i][j] : 0.0;
                }
            __syncthreads();
            PROF(prof, k, 3);
            if (cta_chol64(D, rinv, s_fail, tid)) ok = false;
            PROF(prof, k, 4);
            publish_factor(T.Dd + (size_t)(k + 1) * kCholDdStride, D, rinv, !ok, tid);
            ctr_arrive(cF);
            PROF(prof, k, 5);
        }
        if (worker) {
            // ---------------- U: trailing update over 64x64 tiles of the lower triangle of the step's rows ----------------
            // tiles (ti >= tl) of the active row tiles, then one tile per tl for the right-hand side row
            ctr_wait(cB, done);
            PROF(prof && gb == 1, k, 11);
            double* const Pi = D;
            double* const Pl = T2;
            for (int t = (lead_tile ? 1 : 0) + w; t < ntile; t += nW) {
                const bool rhs = t >= ntri;
                int ti = 0, tl = t - ntri;
                if (!rhs) {
                    ti = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
                    while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
                    while (ti * (ti + 1) / 2 > t) --ti;
                    tl = t - ti * (ti + 1) / 2;
                }
                const int i0 = rhs ? nrows : (ti < n1t ? base + kCB * ti : r2b + kCB * (ti - n1t));
                const int l0 = tl < n1t ? base + kCB * tl : r2b + kCB * (tl - n1t);
                const int ni = rhs ? 1 : kCB; // valid rows of the tile
                __syncthreads();
                const bool diag_tile = !rhs && ti == tl;
                { // all 32 loads of the two panel tiles in flight at once (L2 latency is the cost of a tile, not its bytes)
                    double vi[16], vl[16];
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const int idx = tid + q * kCT, r = idx >> 6, c = idx & 63;
                        vi[q] = r < ni ? __ldcg(A + (size_t)(i0 + r) * ld + j0 + c) : 0.0;
                        vl[q] = __ldcg(A + (size_t)(l0 + r) * ld + j0 + c);
                    }
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const int idx = tid + q * kCT, r = idx >> 6, c = idx & 63;
                        Pi[r * kCS + c] = vi[q], Pl[r * kCS + c] = vl[q];
                    }
                }
                double old[4][4]; // the entries the tile updates, fetched now: their latency hides under the product
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int r = ty + 16 * i, c = tx + 16 * j;
                        old[i][j] = (r < ni && (!diag_tile || c <= r)) ? __ldcg(A + (size_t)(i0 + r) * ld + l0 + c) : 0.0;
                    }
                __syncthreads();
                double acc[4][4] = {};
#pragma unroll 4
                for (int kk = 0; kk < kCB; ++kk) {
                    double av[4], bv[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) av[i] = Pi[(ty + 16 * i) * kCS + kk];
#pragma unroll
                    for (int i = 0; i < 4; ++i) bv[i] = Pl[(tx + 16 * i) * kCS + kk];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int r = ty + 16 * i, c = tx + 16 * j;
                        if (r < ni && (!diag_tile || c <= r)) A[(size_t)(i0 + r) * ld + l0 + c] = old[i][j] - acc[i][j];
                    }
            }
            PROF(prof && gb == 1, k, 12);
            ctr_arrive(cU);
            if (next) ctr_wait(cU, done); // the next panel reads rows other workers updated
            PROF(prof && gb == 1, k, 13);
        }
    }
    ctr_wait(cU, (unsigned)nsteps * (unsigned)nW); // everyone: the factor is complete
    return ok;
}

// Backward substitution  L^T x = y  of a task's eliminated columns, pipelined over the CTAs of its set with flags.
// Block b of x belongs to CTA (nb - 1 - b) mod G.  Its owner subtracts L[b', b]^T x_b' for every row block b' the
// elimination step b touched (border rows: x given and flagged before the call) as soon as x_b' is published (the L tile
// is already in registers by then), solves the 64x64 triangle with one warp and publishes x_b.  Owners take their
// blocks top down and only ever wait for higher blocks, and all CTAs are resident (cooperative launch): no deadlock.
__device__ void task_backward(const CholTask& T, int gb, int G, double* smd)
{
    double* const D = smd;                          // [kCB][kCS] L_bb
    double* const colb = smd + 3 * kCB * kCS + kCB; // [2][kCB]   y_b
    double* const red = colb + 2 * kCB;             // [4][kCB]   partial sums
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ld = T.ld, nrows = T.nrows;
    const double* const A = T.A;
    const int nb = T.n_elim / kCB;
    const int q = tid >> 6, cc = tid & 63;
    for (int b = nb - 1 - gb; b >= 0; b -= G) {
        const int cb = b * kCB;
        const int4 st = __ldg(T.steps + b);
        { // L_bb (row-major)
            const double* dd = T.Dd + (size_t)b * kCholDdStride;
            for (int idx = tid; idx < kCB * kCB; idx += kCT) D[(idx >> 6) * kCS + (idx & 63)] = __ldcg(dd + idx);
        }
        double* const ysh = colb; // y_b (fetched before the waits)
        if (tid < kCB) ysh[tid] = __ldcg(A + (size_t)nrows * ld + cb + tid);
        double acc = 0.0;
        // the row blocks step b touched, highest first: the border range, then the band range
        const int nb2 = max(st.z - st.y, 0) / kCB, nb1 = max(st.x - (cb + kCB), 0) / kCB;
        for (int i = nb1 + nb2 - 1; i >= 0; --i) {
            const int rb = i >= nb1 ? st.y + kCB * (i - nb1) : cb + kCB + kCB * i;
            double t[16];
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) t[jj] = __ldcg(A + (size_t)(rb + q * 16 + jj) * ld + cb + cc);
            if (tid == 0) {
                const int* f = T.flags + rb / kCB;
                int v;
                do {
                    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                } while (v == 0);
                __threadfence();
            }
            __syncthreads();
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) acc = fma(t[jj], __ldcg(T.x + rb + q * 16 + jj), acc);
        }
        red[q * kCB + cc] = acc;
        __syncthreads();
        if (warp == 0) {
            // reciprocals of the diagonal, one pair per lane: the serial chain below multiplies
            const double ri0 = 1.0 / D[lane * kCS + lane];
            const double ri1 = 1.0 / D[(lane + 32) * kCS + lane + 32];
            double y0 = ysh[lane] - (red[lane] + red[kCB + lane] + red[2 * kCB + lane] + red[3 * kCB + lane]);
            double y1 = ysh[lane + 32] - (red[lane + 32] + red[kCB + lane + 32] + red[2 * kCB + lane + 32] + red[3 * kCB + lane + 32]);
            for (int j = kCB - 1; j >= 32; --j) {
                const double xj = shfl_d(y1, j - 32) * shfl_d(ri1, j - 32);
                if (lane + 32 == j) y1 = xj;
                if (lane + 32 < j) y1 = fma(-xj, D[j * kCS + lane + 32], y1);
                y0 = fma(-xj, D[j * kCS + lane], y0);
            }
            for (int j = 31; j >= 0; --j) {
                const double xj = shfl_d(y0, j) * shfl_d(ri0, j);
                if (lane == j) y0 = xj;
                if (lane < j) y0 = fma(-xj, D[j * kCS + lane], y0);
            }
            T.x[cb + lane] = y0;
            T.x[cb + lane + 32] = y1;
            __threadfence();
            __syncwarp();
            if (lane == 0) asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(T.flags + b), "r"(1) : "memory");
        }
        __syncthreads();
    }
}

} // namespace

// zero the envelope of every task matrix (row r: columns 64 * first64[r / 64] .. r; the right-hand side row: all), put the
// identity on the padding columns, and reset flags, barrier counters and the failure flag
__global__ void __launch_bounds__(256) k_dense_zero(LargeSolve L)
{
    int t = 0;
    const int g = blockIdx.x;
    while (g >= L.row_ptr[t + 1]) ++t;
    const CholTask& T = L.tasks[t];
    const int r = g - L.row_ptr[t];
    double* row = T.A + (size_t)r * T.ld;
    if (r == T.nrows) {
        for (int c = threadIdx.x; c < T.nrows; c += blockDim.x) row[c] = 0.0;
        for (int c = threadIdx.x; c < T.nrows / kCB; c += blockDim.x) T.flags[c] = 0;
        if (threadIdx.x < kBarsPerTask) L.bars[kBarsPerTask * t + threadIdx.x] = 0u;
        if (t == 0 && threadIdx.x == 0) *L.fail = 0;
    } else {
        const int c0 = kCB * T.first64[r / kCB];
        const bool pad = r >= T.n_real && r < T.n_elim;
        for (int c = c0 + threadIdx.x; c <= r; c += blockDim.x) row[c] = (pad && c == r) ? 1.0 : 0.0;
    }
}

// one warp per stored S block: scatter the packed 6x6 blocks (and the rhs) into the task matrices.  Block (I, J), I >= J in
// free-pose order: both poses in the same task -> its lower triangle; one of them in a separator -> the border rows of the
// OTHER pose's segment (transposed when the separator pose is J: the separator precedes the segment in the chain).
__global__ void __launch_bounds__(256) k_dense_assemble(DevProblem P, LargeSolve L)
{
    const int lane = threadIdx.x & 31;
    const int blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (blk < P.n_blocks) {
        const int2 ij = P.blk_ij[blk];
        const int tI = L.cam_task[ij.x], tJ = L.cam_task[ij.y];
        const int oI = L.cam_off[ij.x], oJ = L.cam_off[ij.y];
        const CholTask* T;
        int r0, c0;
        bool transpose = false;
        if (tI == tJ) {
            T = L.tasks + tI, r0 = oI, c0 = oJ;
        } else if (tI == L.n_seg) { // I in separator s, J in segment s: the separator is the segment's RIGHT border
            T = L.tasks + tJ, r0 = T->n_elim + T->sL + (oI - L.sep_cols * L.cam_sep[ij.x]), c0 = oJ;
        } else { // J in separator s, I in segment s + 1: the separator is the segment's LEFT border
            T = L.tasks + tI, r0 = T->n_elim + (oJ - L.sep_cols * L.cam_sep[ij.y]), c0 = oI, transpose = true;
        }
        for (int e = lane; e < 36; e += 32) {
            const int r = e / 6, c = e - 6 * r;
            const double v = P.S_blk[36 * (size_t)blk + e];
            if (transpose) T->A[(size_t)(r0 + c) * T->ld + c0 + r] = v;
            else T->A[(size_t)(r0 + r) * T->ld + c0 + c] = v;
        }
    } else if (blk < P.n_blocks + P.n_cams) {
        const int c = blk - P.n_blocks;
        const int f = P.cam_free[c];
        if (f >= 0 && lane < 6) {
            const CholTask& T = L.tasks[L.cam_task[f]];
            T.A[(size_t)T.nrows * T.ld + L.cam_off[f] + lane] = P.bs[6 * (size_t)c + lane];
        }
    }
}

__global__ void __launch_bounds__(kCT, 1) k_chol_coop(LargeSolve L)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) double smd[];
    __shared__ int s_fail;
    const int tid = threadIdx.x;
    const int G = gridDim.x, bid = blockIdx.x;
    const int P = L.n_seg;
    if (tid == 0) s_fail = 0;
    __syncthreads();
    const CholTask& top = L.tasks[P];
    int mine = -1; // the segment this CTA works on in phases 1 and 4
    for (int t = 0; t < P; ++t)
        if (bid >= L.tasks[t].cta0 && bid < L.tasks[t].cta0 + L.tasks[t].ncta) mine = t;
    bool ok = true;
    if (P > 0) {
        // ---------------- phase 1: the interior segments, each on its own CTA group ----------------
        if (mine >= 0) {
            const CholTask& T = L.tasks[mine];
            ok = task_factor(T, bid - T.cta0, T.ncta, L.bars + kBarsPerTask * mine, smd, &s_fail, false);
        }
        grid.sync();
        // ---------------- phase 2: border blocks of the segments -> separator system, in a fixed order ----------------
        const int s = L.sep_cols;
        const long long per = (long long)s * s, tot = (long long)(P - 1) * (2 * per + s);
        for (long long idx = (long long)bid * kCT + tid; idx < tot; idx += (long long)G * kCT) {
            const int j = (int)(idx / (2 * per + s));
            const long long rem = idx - (long long)j * (2 * per + s);
            const CholTask& Ta = L.tasks[j];     // separator j is the RIGHT border of segment j ...
            const CholTask& Tb = L.tasks[j + 1]; // ... and the LEFT border of segment j + 1
            const int ra = Ta.n_elim + Ta.sL, rbb = Tb.n_elim;
            if (rem < per) { // diagonal block of separator j (lower triangle)
                const int r = (int)(rem / s), c = (int)(rem - (long long)r * s);
                if (c <= r) {
                    double* e = top.A + (size_t)(s * j + r) * top.ld + s * j + c;
                    *e = (*e + __ldcg(Ta.A + (size_t)(ra + r) * Ta.ld + ra + c)) + __ldcg(Tb.A + (size_t)(rbb + r) * Tb.ld + rbb + c);
                }
            } else if (rem < 2 * per) { // coupling of separator j with separator j - 1: fill of segment j
                if (j > 0) {
                    const long long q = rem - per;
                    const int r = (int)(q / s), c = (int)(q - (long long)r * s);
                    double* e = top.A + (size_t)(s * j + r) * top.ld + s * (j - 1) + c;
                    *e += __ldcg(Ta.A + (size_t)(ra + r) * Ta.ld + Ta.n_elim + c);
                }
            } else { // right-hand side
                const int c = (int)(rem - 2 * per);
                double* e = top.A + (size_t)top.nrows * top.ld + s * j + c;
                *e = (*e + __ldcg(Ta.A + (size_t)Ta.nrows * Ta.ld + ra + c)) + __ldcg(Tb.A + (size_t)Tb.nrows * Tb.ld + rbb + c);
            }
        }
        grid.sync();
    }
    // ---------------- phase 3: T (the separator system, or the whole matrix) on every CTA ----------------
    {
        const bool okT = task_factor(top, bid, G, L.bars + kBarsPerTask * P, smd, &s_fail, true);
        ok = ok && okT;
        PROF(bid == 0, top.n_elim / kCB, 0);
        task_backward(top, bid, G, smd);
        PROF(bid == 0, top.n_elim / kCB, 1);
    }
    if (!ok && tid == 0) *L.fail = 1;
    if (P > 0) {
        grid.sync();
        // ---------------- phase 4: separator values into the borders, then the interiors ----------------
        if (mine >= 0) {
            const CholTask& T = L.tasks[mine];
            const int gb = bid - T.cta0;
            const int s = L.sep_cols;
            for (int i = gb * kCT + tid; i < T.sL + T.sR; i += T.ncta * kCT) {
                const int sep = i < T.sL ? mine - 1 : mine;
                T.x[T.n_elim + i] = __ldcg(top.x + s * sep + (i < T.sL ? i : i - T.sL));
            }
            for (int i = gb * kCT + tid; i < (T.sL + T.sR) / kCB; i += T.ncta * kCT) T.flags[T.n_elim / kCB + i] = 1;
            group_barrier(L.bars + kBarsPerTask * mine + 3, (unsigned)T.ncta);
            task_backward(T, gb, T.ncta, smd);
        }
    }
}

__global__ void __launch_bounds__(256) k_large_finish(DevProblem P, LargeSolve L)
{
    __shared__ double sh[256];
    LmState& st = P.lm[0];
    if (!st.active) return;
    const int fail = *L.fail;
    const int cur = st.cur, cand = cur ^ 1;
    const int tid = threadIdx.x;
    double ss = 0.0;
    for (int c = tid; c < P.n_cams; c += blockDim.x) {
        const int f = P.cam_free[c];
        double d[6];
        for (int r = 0; r < 6; ++r) d[r] = 0.0;
        if (f >= 0 && !fail) {
            const double* x = L.tasks[L.cam_task[f]].x + L.cam_off[f];
            for (int r = 0; r < 6; ++r) d[r] = x[r], L.x[6 * f + r] = x[r];
        }
        for (int r = 0; r < 6; ++r) P.dp[6 * (size_t)c + r] = d[r], ss += d[r] * d[r];
        const double* pc = P.pose[cur] + 7 * (size_t)c;
        double* pn = P.pose[cand] + 7 * (size_t)c;
        if (f >= 0 && !fail) pose_oplus(d, pc, pn);
        else
            for (int r = 0; r < 7; ++r) pn[r] = pc[r];
        write_cam_rt(pn, P.cam_rt[cand] + 12 * (size_t)c);
    }
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

int large_solve_max_n() { return 6 * 5000; } // unpartitioned: 7.2 GB of dense lower triangle; nothing else limits n

int large_solve_cta_count()
{
    static PerDeviceOnce once;
    return once.sm_count();
}

// the factorisation + both substitutions alone (matrices assembled, x out); cooperative launch: the runtime guarantees
// that all CTAs are co-resident, which the grid and group barriers need
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

int launch_dense_zero(const LargeSolve& L, cudaStream_t st)
{
    k_dense_zero<<<L.total_rows, 256, 0, st>>>(L);
    return 1;
}

int launch_solve_large(const DevProblem& P, const LargeSolve& L, cudaStream_t st)
{
    const long long warps = (long long)P.n_blocks + P.n_cams;
    k_dense_zero<<<L.total_rows, 256, 0, st>>>(L);
    k_dense_assemble<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(P, L);
    const int k = launch_chol_coop(L, st);
    k_large_finish<<<1, 256, 0, st>>>(P, L);
    return 3 + k;
}

} // namespace visoba
