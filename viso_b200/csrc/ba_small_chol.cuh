// ba_small_chol.cuh — dense Cholesky solve of ONE small reduced camera system (n <= kSmallMaxN) by one CTA in shared
// memory: g2o LinearSolverDense on the pose block of BlockSolver<6,3> (reference include/bundle_adjuster.h:113-114).
// Shared by k_solve_small (one CTA per problem of a batch, ba_kernels.cu) and the single-window LM kernel (ba_window.cu).
//
// Blocked right-looking factorisation, 8 columns per step (the column-at-a-time version it replaces spent three CTA
// barriers and a shared-memory round trip per column):
//     - the 8 x 8 diagonal block by ONE thread in registers: the pivot chain (rsqrt + two dependent FMAs per pivot);
//     - panel: one thread per row below it,  x L88^T = a;
//     - remainder -= panel panel^T in 2 x 2 micro-tiles of the lower triangle by warps 1..; warp 0 (whose lanes 0..7 own the
//       panel rows of the NEXT diagonal block) arrives at the panel barrier without waiting, updates that block and its
//       lane 0 factorises it at once: the pivot chain runs beside the remainder and never waits for a full-CTA barrier.
// The right-hand side rides along as one more row, so the forward substitution falls out; the backward substitution
// is one warp with the solution in registers (shuffle -> multiply -> FMA per unknown).  Everything is fp64 and in a fixed
// order: bit-identical from run to run and between the two callers.
#ifndef VISO_BA_SMALL_CHOL_CUH
#define VISO_BA_SMALL_CHOL_CUH

#include "ba_device.h"

namespace visoba {

constexpr int kSmallN8 = (kSmallMaxN + 7) & ~7; // 168

__host__ __device__ constexpr int small_ld(int n8) { return (n8 + 1) | 1; } // odd row stride: column walks are conflict-free
__host__ __device__ constexpr size_t small_smem_bytes(int n8)
{ // matrix [n8 + 1][ld] (row n8: the right-hand side), then 1 / diag
    return sizeof(double) * ((size_t)(n8 + 1) * (size_t)small_ld(n8) + (size_t)n8 + 8);
}

__device__ __forceinline__ void small_chol8(double* A, int ld, double* rinv, int c0, int* s_fail)
{
    double a[36];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) a[i * (i + 1) / 2 + j] = A[(c0 + i) * ld + c0 + j];
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
        for (int j = 0; j <= i; ++j) A[(c0 + i) * ld + c0 + j] = a[i * (i + 1) / 2 + j];
}

// In place: L of the leading n8 x n8 lower triangle (n8 a multiple of 8, padded with the identity), and row n8 (the
// right-hand side b) becomes y = L^-1 b.  nt threads (a multiple of 32, >= 64).  *s_fail is set on a bad pivot.
__device__ __forceinline__ void small_chol_forward(double* A, int n8, int ld, double* rinv, int* s_fail, int tid, int nt)
{
    if (tid == 0) small_chol8(A, ld, rinv, 0, s_fail);
    __syncthreads();
#pragma unroll 1
    for (int c0 = 0; c0 < n8; c0 += 8) {
        const int mt = n8 - c0 - 8; // rows of the triangle below the diagonal block; one more row: the right-hand side
        for (int r = tid; r <= mt; r += nt) { // panel
            double* row = A + (c0 + 8 + r) * ld + c0;
            double x[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = row[j];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                double s = x[j];
#pragma unroll
                for (int t = 0; t < j; ++t) s = fma(-x[t], A[(c0 + j) * ld + c0 + t], s);
                x[j] = s * rinv[c0 + j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) row[j] = x[j];
        }
        if (mt == 0) break;
        if (tid < 32) { // the next diagonal block (36 entries: lanes 0..3 take two), then its factorisation by lane 0
            __syncwarp();  // its panel rows are rows 0..7 of this warp: no need to wait for the other warps' rows
            __threadfence_block();
            asm volatile("bar.arrive 1, %0;" ::"r"(nt) : "memory");
            // entry e = tid of the packed lower triangle: row rr, column cc (no loop: this is the pivot chain's time)
            const int rr = tid >= 28 ? 7 : tid >= 21 ? 6 : tid >= 15 ? 5 : tid >= 10 ? 4 : tid >= 6 ? 3 : tid >= 3 ? 2 : tid >= 1 ? 1 : 0;
            const int cc = tid - rr * (rr + 1) / 2; // tid = 28..31 are the first entries of row 7; entries 32..35 the last four
            const bool two = tid < 4;
            const double* pr = A + (c0 + 8 + rr) * ld + c0;
            const double* pc = A + (c0 + 8 + cc) * ld + c0;
            const double* pr2 = A + (c0 + 8 + 7) * ld + c0;
            const double* pc2 = A + (c0 + 8 + 4 + (tid & 3)) * ld + c0;
            // one 8-term sum per entry, in the order every other entry of the remainder uses (same arithmetic for all)
            double s = 0.0, s2 = 0.0;
#pragma unroll
            for (int t = 0; t < 8; ++t) s = fma(pr[t], pc[t], s), s2 = fma(pr2[t], pc2[t], s2);
            A[(c0 + 8 + rr) * ld + c0 + 8 + cc] -= s;
            if (two) A[(c0 + 8 + 7) * ld + c0 + 8 + 4 + tid] -= s2;
            __syncwarp();
            if (tid == 0) small_chol8(A, ld, rinv, c0 + 8, s_fail);
        } else {
            asm volatile("bar.sync 1, %0;" ::"r"(nt) : "memory"); // every panel row is written
            // triangle below the next diagonal block in 2 x 2 micro-tiles: tile rows R = 4 .. h - 1 hold R + 1 tiles each; row
            // 4 + rp is folded with row h - 1 - rp into one line of W = h + 5 tiles (rectangular enumeration, no idle
            // lanes on the upper triangle); then the right-hand side row, one entry per work item
            const int h = mt >> 1, W = h + 5, nline = (h - 4) >> 1, ntile = nline * W;
            for (int idx = tid - 32; idx < ntile + mt; idx += nt - 32) {
                if (idx < ntile) {
                    const int rp = idx / W, q = idx - rp * W;
                    int R = 4 + rp, C = q;
                    if (q > R) R = h - 1 - rp, C = q - (5 + rp);
                    const double* pr0 = A + (c0 + 8 + 2 * R) * ld + c0;
                    const double* pc0 = A + (c0 + 8 + 2 * C) * ld + c0;
                    double s00 = 0.0, s01 = 0.0, s10 = 0.0, s11 = 0.0;
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        const double r0 = pr0[t], r1 = pr0[ld + t], q0 = pc0[t], q1 = pc0[ld + t];
                        s00 = fma(r0, q0, s00), s01 = fma(r0, q1, s01), s10 = fma(r1, q0, s10), s11 = fma(r1, q1, s11);
                    }
                    double* o = A + (c0 + 8 + 2 * R) * ld + c0 + 8 + 2 * C;
                    o[0] -= s00;
                    if (C < R) o[1] -= s01;
                    o[ld] -= s10, o[ld + 1] -= s11;
                } else { // right-hand side row (row mt of the remainder), column cc of the remainder
                    const int cc = idx - ntile;
                    if (cc < 8) continue; // the next diagonal block's columns: below
                    const double* pr = A + (c0 + 8 + mt) * ld + c0;
                    const double* pc = A + (c0 + 8 + cc) * ld + c0;
                    double s = 0.0;
#pragma unroll
                    for (int t = 0; t < 8; ++t) s = fma(pr[t], pc[t], s);
                    A[(c0 + 8 + mt) * ld + c0 + 8 + cc] -= s;
                }
            }
            // the right-hand side entries under the next diagonal block: 8 items, by the first lanes of warp 1
            if (tid >= 32 && tid < 40) {
                const int cc = tid - 32;
                const double* pr = A + (c0 + 8 + mt) * ld + c0;
                const double* pc = A + (c0 + 8 + cc) * ld + c0;
                double s = 0.0;
#pragma unroll
                for (int t = 0; t < 8; ++t) s = fma(pr[t], pc[t], s);
                A[(c0 + 8 + mt) * ld + c0 + 8 + cc] -= s;
            }
        }
        __syncthreads();
    }
}

// L^T x = y by warp 0 (call with the whole warp; y = row n8 of A, overwritten by x).  x in registers: lane l holds
// entries l, l + 32, ...; per unknown: shuffle -> multiply by 1 / diag -> FMA into the entries above.
__device__ __forceinline__ void small_backward_warp(double* A, int n8, int ld, const double* rinv, int lane)
{
    constexpr int kSeg = (kSmallN8 + 31) / 32; // 6
    double* y = A + n8 * ld;
    double v[kSeg];
#pragma unroll
    for (int s = 0; s < kSeg; ++s) v[s] = (32 * s + lane < n8) ? y[32 * s + lane] : 0.0;
#pragma unroll
    for (int s = kSeg - 1; s >= 0; --s) {
        if (32 * s >= n8) continue;
        const int jtop = min(n8 - 1, 32 * s + 31);
        for (int j = jtop; j >= 32 * s; --j) {
            const double xj = __shfl_sync(0xffffffffu, v[s], j - 32 * s) * rinv[j];
            const double* Lj = A + j * ld; // row j of L = column j of L^T
            if (lane + 32 * s == j) v[s] = xj;
            if (lane + 32 * s < j) v[s] = fma(-xj, Lj[32 * s + lane], v[s]);
#pragma unroll
            for (int t = 0; t < kSeg; ++t)
                if (t < s) v[t] = fma(-xj, Lj[32 * t + lane], v[t]);
        }
    }
#pragma unroll
    for (int s = 0; s < kSeg; ++s)
        if (32 * s + lane < n8) y[32 * s + lane] = v[s];
}

} // namespace visoba
#endif
