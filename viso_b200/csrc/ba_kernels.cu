// ba_kernels.cu — hand-written sm_100a kernels of the viso bundle-adjustment hot path.
//
// Reference semantics restated on the device (citations relative to /root/reference):
//   projection          include/keyframe.h:82-89          Pc = R Pw + t, u = fx x/z + cx
//   2x6 pose Jacobian   src/viso.cpp:640-658 (dPixeldXi)  columns (v, w), left perturbation
//   2x3 point Jacobian  first three columns of dPixeldXi times R (absent in the reference)
//   pose update         include/bundle_adjuster.h:51-55   T <- exp(d) T
//   point update        include/bundle_adjuster.h:34-38   P <- P + d
//   6/3 block Schur     include/bundle_adjuster.h:108,:128 (g2o BlockSolver<6,3>, points marginalised)
//   dense pose solve    include/bundle_adjuster.h:113-114 (g2o LinearSolverDense)
//   LM                  include/bundle_adjuster.h:116-117 (g2o OptimizationAlgorithmLevenberg)
//
// No tensor cores: the blocks are 6x6 / 3x3 fp64.  Everything is scatter-free: per-point
// quantities are owned by one thread, camera blocks are GATHERED through a pair list that
// is sorted by destination block, and every cross-thread sum is a fixed-shape tree, so the
// result is bit-reproducible and independent of the order edges were added in.
#include "ba_math.cuh"
#include "ba_small_chol.cuh"

namespace visoba {

// ------------------------------------------------------------------------------------------
// state init / R|t cache
// ------------------------------------------------------------------------------------------
__global__ void k_state_init(DevProblem P, int iterations, int max_trials, double lambda_init)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0) *P.n_active = P.n_problems;
    if (b >= P.n_problems) return;
    LmState& s = P.lm[b];
    s.lambda = lambda_init;
    s.nu = 2.0;
    s.chi = 0.0;
    s.chi_start = 0.0;
    s.chi_initial = 0.0;
    s.rho = 0.0;
    s.dp_sq = 0.0;
    s.it = 0;
    s.q = 0;
    s.trials = 0;
    s.active = iterations > 0 ? 1 : 0;
    s.terminated = 0;
    s.ok2 = 1;
    s.accepted = 0;
    s.target_it = iterations;
    s.max_trials = max_trials;
    P.maxdiag_bits[b] = 0ull;
    if (iterations <= 0 && b == 0) *P.n_active = 0;
}

__global__ void k_cam_rt(DevProblem P, int which)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.n_cams) return;
    if (which < 0) which = P.lm[P.cam_prob[c]].cur; // the current buffer of the camera's problem
    write_cam_rt(P.pose[which] + 7 * (size_t)c, P.cam_rt[which] + 12 * (size_t)c);
}

// read-back: current poses / points of every problem into one contiguous buffer each
__global__ void k_gather_current(DevProblem P, double* __restrict__ poses_out, double* __restrict__ pts_out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n_pose = 7ll * P.n_cams, n_pt = 3ll * P.n_points;
    if (i < n_pose) {
        const int c = (int)(i / 7);
        poses_out[i] = P.pose[P.lm[P.cam_prob[c]].cur][i];
    } else if (i < n_pose + n_pt) {
        const long long j = i - n_pose;
        const int p = (int)(j / 3);
        pts_out[j] = P.pts[P.lm[P.pt_prob[p]].cur][j];
    }
}

// ------------------------------------------------------------------------------------------
// K1 linearize: one thread per point.
//   pass 1  H_ll = sum w Jl^T Jl, b_l = -sum w Jl^T r, robust cost of the point
//   3x3     L L^T = H_ll + lambda I, L^-1, zb = L^-1 b_l          -> ptinfo (72 B)
//   pass 2  per observation record {x/z, y/z, 1/z, w, G~ = F w Jl L^-T, v~ = F (w r + G zb)}
// The record is all the camera-block gather (K2) and the back substitution (K4) need:
//   S_IJ  -= Jp_a^T G_a G_b^T Jp_b,  S_II += w Jp_a^T Jp_a,  bs_I -= Jp_a^T v_a
// with Jp = F J~(x/z, y/z, 1/z) rebuilt from three doubles (viso.cpp:653-655).
// ------------------------------------------------------------------------------------------
// R|t of one pose (96 bytes, 16-byte aligned) by six 16-byte loads instead of twelve 8-byte ones: the per-point kernels
// below are bound by the number of their outstanding global loads (ncu on C4: 56 % long scoreboard, 29 % LG throttle)
__device__ __forceinline__ void load_rt(const double* __restrict__ cam_rt, int cam, double (&rt)[12])
{
    const double2* __restrict__ p = reinterpret_cast<const double2*>(cam_rt + 12 * (size_t)cam);
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const double2 v = __ldg(p + k);
        rt[2 * k] = v.x, rt[2 * k + 1] = v.y;
    }
}

// 32-byte global accesses (sm_100: LDG / STG.256).  A per-observation record is three 32-byte sectors; written as six 16-byte
// pieces by lanes whose records lie 1 KB apart, every store instruction became 32 half-sector requests to L2, and the
// request rate, not the bytes, bounded k_linearize.  p must be 32-byte aligned.
__device__ __forceinline__ void st_global_256(double* p, double a, double b, double c, double d)
{
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
__device__ __forceinline__ void ld_global_256(const double* p, double& a, double& b, double& c, double& d)
{
    asm volatile("ld.global.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}

__global__ void __launch_bounds__(128) k_linearize(DevProblem P, int init_pass, double lambda_override)
{
    const int p = P.pt_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.pt_end) return;
    const int prob = P.pt_prob[p];
    const LmState& st = P.lm[prob];
    if (!st.active && lambda_override < 0.0) return;
    const int cur = st.cur;
    const double lambda = init_pass ? 0.0 : (lambda_override >= 0.0 ? lambda_override : st.lambda);
    const double* __restrict__ pt = P.pts[cur] + 3 * (size_t)p;
    const double X = pt[0], Y = pt[1], Z = pt[2];
    const double* __restrict__ cam_rt = P.cam_rt[cur];
    const int s0 = P.pt_ptr[p], s1 = P.pt_ptr[p + 1];

    double h00 = 0, h10 = 0, h11 = 0, h20 = 0, h21 = 0, h22 = 0, b0 = 0, b1 = 0, b2 = 0, cost = 0;
    // the pose index and the pixel of the NEXT observation are requested before this one's R|t: one dependent round trip less
    int cam_n = s0 < s1 ? __ldg(P.obs_cam + s0) : 0;
    float2 uv_n = s0 < s1 ? __ldg(P.obs_uv + s0) : make_float2(0.f, 0.f);
    for (int s = s0; s < s1; ++s) {
        const int cam = cam_n;
        const float2 uv = uv_n;
        if (s + 1 < s1) cam_n = __ldg(P.obs_cam + s + 1), uv_n = __ldg(P.obs_uv + s + 1);
        double rt[12];
        load_rt(cam_rt, cam, rt);
        ObsGeom g;
        obs_geometry(P, rt, X, Y, Z, uv, g);
        cost += g.rho;
        h00 += g.w * (g.jl0[0] * g.jl0[0] + g.jl1[0] * g.jl1[0]);
        h10 += g.w * (g.jl0[1] * g.jl0[0] + g.jl1[1] * g.jl1[0]);
        h11 += g.w * (g.jl0[1] * g.jl0[1] + g.jl1[1] * g.jl1[1]);
        h20 += g.w * (g.jl0[2] * g.jl0[0] + g.jl1[2] * g.jl1[0]);
        h21 += g.w * (g.jl0[2] * g.jl0[1] + g.jl1[2] * g.jl1[1]);
        h22 += g.w * (g.jl0[2] * g.jl0[2] + g.jl1[2] * g.jl1[2]);
        b0 -= g.w * (g.jl0[0] * g.ru + g.jl1[0] * g.rv);
        b1 -= g.w * (g.jl0[1] * g.ru + g.jl1[1] * g.rv);
        b2 -= g.w * (g.jl0[2] * g.ru + g.jl1[2] * g.rv);
    }
    P.cost_pt[p] = cost;
    double i00 = 0, i10 = 0, i11 = 0, i20 = 0, i21 = 0, i22 = 0, z0 = 0, z1 = 0, z2 = 0;
    if (init_pass) {
        // g2o computeLambdaInit: the point part of max diag(H); max is order-free -> deterministic
        double m = fmax(h00, fmax(h11, h22));
        // a full warp of points of one problem (the rule: one large problem has 100k points on ONE address) folds
        // its maximum first and sends one atomic; ragged warps keep one atomic per point
        const unsigned act = __activemask();
        const bool whole = act == 0xffffffffu && __match_any_sync(act, prob) == 0xffffffffu;
        if (whole) {
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, off));
        }
        if (m > 0.0 && (!whole || (threadIdx.x & 31) == 0))
            atomicMax(P.maxdiag_bits + prob, (unsigned long long)__double_as_longlong(m));
    } else {
        h00 += lambda;
        h11 += lambda;
        h22 += lambda;
        const double l00 = sqrt(h00);
        i00 = 1.0 / l00;
        const double l10 = h10 * i00, l20 = h20 * i00;
        const double l11 = sqrt(h11 - l10 * l10);
        i11 = 1.0 / l11;
        const double l21 = (h21 - l20 * l10) * i11;
        const double l22 = sqrt(h22 - l20 * l20 - l21 * l21);
        i22 = 1.0 / l22;
        i10 = -l10 * i00 * i11;
        i21 = -l21 * i11 * i22;
        i20 = -(l20 * i00 + l21 * i10) * i22;
        z0 = i00 * b0;
        z1 = i10 * b0 + i11 * b1;
        z2 = i20 * b0 + i21 * b1 + i22 * b2;
    }
    double* info = P.ptinfo + kPtInfoDoubles * (size_t)p;
    info[0] = i00, info[1] = i10, info[2] = i11, info[3] = i20, info[4] = i21, info[5] = i22;
    info[6] = z0, info[7] = z1, info[8] = z2;

    cam_n = s0 < s1 ? __ldg(P.obs_cam + s0) : 0;
    uv_n = s0 < s1 ? __ldg(P.obs_uv + s0) : make_float2(0.f, 0.f);
    for (int s = s0; s < s1; ++s) {
        const int cam = cam_n;
        const float2 uv = uv_n;
        if (s + 1 < s1) cam_n = __ldg(P.obs_cam + s + 1), uv_n = __ldg(P.obs_uv + s + 1);
        double rt[12];
        load_rt(cam_rt, cam, rt);
        ObsGeom g;
        obs_geometry(P, rt, X, Y, Z, uv, g);
        // G = w Jl L^-T   (L^-T)[m][k] = Linv[k][m], k >= m
        const double g00 = g.w * (g.jl0[0] * i00);
        const double g01 = g.w * (g.jl0[0] * i10 + g.jl0[1] * i11);
        const double g02 = g.w * (g.jl0[0] * i20 + g.jl0[1] * i21 + g.jl0[2] * i22);
        const double g10 = g.w * (g.jl1[0] * i00);
        const double g11 = g.w * (g.jl1[0] * i10 + g.jl1[1] * i11);
        const double g12 = g.w * (g.jl1[0] * i20 + g.jl1[1] * i21 + g.jl1[2] * i22);
        const double v0 = g.w * g.ru + g00 * z0 + g01 * z1 + g02 * z2;
        const double v1 = g.w * g.rv + g10 * z0 + g11 * z1 + g12 * z2;
        double* r = P.rec + kRecDoubles * (size_t)s; // 96 bytes per record, the array 256-byte aligned
        st_global_256(r, g.xn, g.yn, g.iz, g.w);
        st_global_256(r + 4, P.fx * g00, P.fx * g01, P.fx * g02, P.fy * g10);
        st_global_256(r + 8, P.fy * g11, P.fy * g12, P.fx * v0, P.fy * v1);
    }
}

// cost of the current state only (viso_ba_eval_cost): same pass 1 without the rest
__global__ void __launch_bounds__(128) k_eval_cost(DevProblem P)
{
    const int p = P.pt_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.pt_end) return;
    const int cur = P.lm[P.pt_prob[p]].cur;
    const double* __restrict__ pt = P.pts[cur] + 3 * (size_t)p;
    const double X = pt[0], Y = pt[1], Z = pt[2];
    const double* __restrict__ cam_rt = P.cam_rt[cur];
    double cost = 0;
    for (int s = P.pt_ptr[p]; s < P.pt_ptr[p + 1]; ++s) {
        ObsGeom g;
        obs_geometry(P, cam_rt + 12 * (size_t)__ldg(P.obs_cam + s), X, Y, Z, __ldg(P.obs_uv + s), g);
        cost += g.rho;
    }
    P.cost_pt[p] = cost;
}

// ------------------------------------------------------------------------------------------
// K2 schur_blocks: one warp per work item (a slice of the pair list of one S block).
// Each lane gathers two 96-byte records per pair, forms  X = J~_a^T M J~_b  (6x6) with
// M = -G~_a G~_b^T (+ diag(w fx^2, w fy^2) when a == b) and keeps 36 + 6 fp64 accumulators in
// registers.  The warp then folds them with a recursive-halving butterfly (31 shuffles for
// 32 values instead of 160) and lane e writes entry e of the item's partial block.
// ------------------------------------------------------------------------------------------
template <int N>
__device__ __forceinline__ void butterfly_reduce(double (&v)[N], int lane)
{ // N = 32: afterwards lane l holds the warp total of v[l] in v[0]
#pragma unroll
    for (int off = 16, cnt = N / 2; off >= 1; off >>= 1, cnt >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < cnt; ++i) {
            const double send = upper ? v[i] : v[i + cnt];
            const double keep = upper ? v[i + cnt] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
}

// Round 2: the records of the next kSbStages pairs of a lane are on their way to shared memory (cp.async, no registers
// held) while the lane computes: the kernel was bound by the latency of its gathers (two dependent L2 round trips per
// pair at 8 warps per SM), not by their bandwidth.  The pair product is the tile kernel's (ba_window.cu): M = -G~_a G~_b^T,
// T = M J~_b, S += J~_a^T T with the structural zeros of J~ left out, fused into the 42 accumulators (104 FP64 instructions
// per pair; the version it replaces formed the 6 x 6 product first and added it: 160).
constexpr int kSbThreads = 128;
constexpr int kSbStages = 4;    // pairs of a lane in flight
constexpr int kSbChunks = 11;   // 16-byte chunks staged per pair: a0..a5, b0..b4
constexpr int kSbMaxIter = kItemMaxPairs / 32; // pairs of an item per lane (general_items_and_chunks cuts the items)
constexpr size_t kSbIdxBytes = sizeof(int2) * kSbMaxIter * kSbThreads;
constexpr size_t kSbBytes = kSbIdxBytes + sizeof(double2) * kSbStages * kSbChunks * kSbThreads;

__device__ __forceinline__ void sb_cp16(void* smem, const void* gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void sb_cp8(void* smem, const void* gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void sb_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void sb_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// v += J~_a^T (M J~_b) for one pair; a*, b*: chunks 0..4 of the two records
__device__ __forceinline__ void sb_pair_product(double (&v)[48], const double2& a0, const double2& a1, const double2& b0, const double2& b1,
    double m00, double m01, double m10, double m11)
{
    const double xb = b0.x, yb = b0.y, zb = b1.x;
    const double bxz = xb * zb, byz = yb * zb, bxy = xb * yb, bxx = fma(xb, xb, 1.0), byy = fma(yb, yb, 1.0);
    double t0[6], t1[6];
    t0[0] = m00 * zb, t1[0] = m10 * zb;
    t0[1] = m01 * zb, t1[1] = m11 * zb;
    t0[2] = -fma(m00, bxz, m01 * byz), t1[2] = -fma(m10, bxz, m11 * byz);
    t0[3] = -fma(m00, bxy, m01 * byy), t1[3] = -fma(m10, bxy, m11 * byy);
    t0[4] = fma(m00, bxx, m01 * bxy), t1[4] = fma(m10, bxx, m11 * bxy);
    t0[5] = fma(m01, xb, -(m00 * yb)), t1[5] = fma(m11, xb, -(m10 * yb));
    const double xa = a0.x, ya = a0.y, za = a1.x;
    const double axz = -(xa * za), ayz = -(ya * za), axy = xa * ya, axx = fma(xa, xa, 1.0), ayy = -fma(ya, ya, 1.0);
    const double naxy = -axy, nya = -ya;
#pragma unroll
    for (int c = 0; c < 6; ++c) {
        v[c] = fma(za, t0[c], v[c]);
        v[6 + c] = fma(za, t1[c], v[6 + c]);
        v[12 + c] = fma(axz, t0[c], fma(ayz, t1[c], v[12 + c]));
        v[18 + c] = fma(naxy, t0[c], fma(ayy, t1[c], v[18 + c]));
        v[24 + c] = fma(axx, t0[c], fma(axy, t1[c], v[24 + c]));
        v[30 + c] = fma(nya, t0[c], fma(xa, t1[c], v[30 + c]));
    }
}

// PHOTO: photometric edges (ba_photo.cu): the 2 x 2 middle of a diagonal pair is the symmetric matrix w F C_tt F in ph_mid
// instead of diag(w fx^2, w fy^2)
template <bool PHOTO>
__global__ void __launch_bounds__(kSbThreads, 2) k_schur_blocks(DevProblem P, int diag_only)
{
    extern __shared__ __align__(16) unsigned char sb_sm[];
    int2* const s_idx = reinterpret_cast<int2*>(sb_sm);                  // [kSbMaxIter][kSbThreads]
    double2* const s_rec = reinterpret_cast<double2*>(sb_sm + kSbIdxBytes); // [kSbStages][kSbChunks][kSbThreads]
    const int tid = threadIdx.x, lane = tid & 31;
    const int widx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int item;
    if (diag_only) {
        if (widx >= P.n_diag_items) return;
        item = P.diag_items[widx];
    } else {
        if (widx >= P.n_items) return;
        item = widx;
    }
    const int4 it = P.items[item];
    const int prob = it.w >> 1;
    const bool is_diag = (it.w & 1) != 0;
    if (!P.lm[prob].active) return;
    const double fx2 = P.fx * P.fx, fy2 = P.fy * P.fy;
    // this lane's pairs: it.x + lane, + 32, ...  Their indices first (one group), then the records of kSbStages pairs ahead
    const int n_it = it.y - it.x > lane ? (it.y - it.x - lane + 31) >> 5 : 0; // <= kSbMaxIter
    for (int j = 0; j < n_it; ++j) sb_cp8(s_idx + j * kSbThreads + tid, P.pairs + it.x + lane + 32 * j);
    sb_commit();
    sb_wait<0>();
    auto issue = [&](int j) { // records of pair j of this lane -> stage j % kSbStages
        if (j < n_it) {
            const int2 pr = s_idx[j * kSbThreads + tid];
            const double2* ra = reinterpret_cast<const double2*>(P.rec + kRecDoubles * (size_t)pr.x);
            double2* dst = s_rec + (size_t)(j % kSbStages) * kSbChunks * kSbThreads + tid;
#pragma unroll
            for (int c = 0; c < 6; ++c) sb_cp16(dst + c * kSbThreads, ra + c);
            if (pr.y != pr.x) {
                const double2* rb = reinterpret_cast<const double2*>(P.rec + kRecDoubles * (size_t)pr.y);
#pragma unroll
                for (int c = 0; c < 5; ++c) sb_cp16(dst + (6 + c) * kSbThreads, rb + c);
            }
        }
        sb_commit(); // one group per pair, empty or not: the wait below counts groups
    };
#pragma unroll
    for (int j = 0; j < kSbStages; ++j) issue(j);

    double v[48]; // 36 block entries, 6 rhs entries, padding
#pragma unroll
    for (int i = 0; i < 48; ++i) v[i] = 0.0;
    for (int j = 0; j < n_it; ++j) {
        sb_wait<kSbStages - 1>(); // the group of pair j has landed
        const int2 pr = s_idx[j * kSbThreads + tid];
        const bool same = pr.x == pr.y;
        const double2* src = s_rec + (size_t)(j % kSbStages) * kSbChunks * kSbThreads + tid;
        const double2 a0 = src[0], a1 = src[kSbThreads], a2 = src[2 * kSbThreads], a3 = src[3 * kSbThreads], a4 = src[4 * kSbThreads];
        const double2 a5 = src[5 * kSbThreads];
        const double2* sb = same ? src : src + 6 * kSbThreads;
        const double2 b0 = sb[0], b1 = sb[kSbThreads], b2 = sb[2 * kSbThreads], b3 = sb[3 * kSbThreads], b4 = sb[4 * kSbThreads];
        issue(j + kSbStages); // the stage is free: its values are in registers
        // M = -G~_a G~_b^T;  G~ = (c2.x c2.y c3.x ; c3.y c4.x c4.y)
        double m00 = -fma(a2.x, b2.x, fma(a2.y, b2.y, a3.x * b3.x));
        double m01 = -fma(a2.x, b3.y, fma(a2.y, b4.x, a3.x * b4.y));
        double m10 = -fma(a3.y, b2.x, fma(a4.x, b2.y, a4.y * b3.x));
        double m11 = -fma(a3.y, b3.y, fma(a4.x, b4.x, a4.y * b4.y));
        if (same) {
            if (PHOTO) {
                const double* __restrict__ md = P.ph_mid + 4 * (size_t)pr.x;
                m00 += md[0], m01 += md[1], m10 += md[1], m11 += md[2];
            } else {
                m00 = fma(a1.y, fx2, m00);
                m11 = fma(a1.y, fy2, m11);
            }
        }
        sb_pair_product(v, a0, a1, b0, b1, m00, m01, m10, m11);
        if (is_diag && !same) // duplicate (point, pose) edges: (a, b) and (b, a) both land in this diagonal block
            sb_pair_product(v, b0, b1, a0, a1, m00, m10, m01, m11);
        if (same) { // rhs: Jp^T v~
            const double xa = a0.x, ya = a0.y, za = a1.x;
            v[36] = fma(za, a5.x, v[36]);
            v[37] = fma(za, a5.y, v[37]);
            v[38] = fma(-(xa * za), a5.x, fma(-(ya * za), a5.y, v[38]));
            v[39] = fma(-(xa * ya), a5.x, fma(-fma(ya, ya, 1.0), a5.y, v[39]));
            v[40] = fma(fma(xa, xa, 1.0), a5.x, fma(xa * ya, a5.y, v[40]));
            v[41] = fma(-ya, a5.x, fma(xa, a5.y, v[41]));
        }
    }
    sb_wait<0>();
    double acc[32], acc2[16]; // the fold below takes 32 + 16 values
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] = v[i];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc2[i] = v[32 + i];
    butterfly_reduce<32>(acc, lane);
    // second group: 16 padded values; fold the two half-warps first, then butterfly over 16 lanes
#pragma unroll
    for (int i = 0; i < 16; ++i) acc2[i] += __shfl_xor_sync(0xffffffffu, acc2[i], 16);
#pragma unroll
    for (int off = 8, cnt = 8; off >= 1; off >>= 1, cnt >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < cnt; ++i) {
            const double send = upper ? acc2[i] : acc2[i + cnt];
            const double keep = upper ? acc2[i + cnt] : acc2[i];
            acc2[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    double* out = P.item_part + kItemPart * (size_t)item;
    out[lane] = acc[0];
    if (lane < 10) out[32 + lane] = acc2[0];
}

// K2b: one warp per S block sums the item partials in list order, adds lambda on the
// diagonal (g2o setLambda) and writes S_blk / bs.  bs sign: g2o solves H x = b, b = -J^T r.
__global__ void __launch_bounds__(128) k_block_finalize(DevProblem P, int init_pass, double lambda_override)
{
    const int lane = threadIdx.x & 31;
    const int blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (blk >= P.n_blocks) return;
    const int cam = P.block_cam[blk];
    if (init_pass && cam < 0) return;
    const int i0 = P.block_item_ptr[blk], i1 = P.block_item_ptr[blk + 1];
    int prob = 0;
    if (cam >= 0) prob = P.cam_prob[cam];
    else if (i0 < i1) prob = P.items[i0].w >> 1;
    else { // block with no co-visible point: keep it zero
        for (int e = lane; e < 36; e += 32) P.S_blk[36 * (size_t)blk + e] = 0.0;
        return;
    }
    const LmState& st = P.lm[prob];
    if (!st.active && lambda_override < 0.0) return;
    const double lambda = init_pass ? 0.0 : (lambda_override >= 0.0 ? lambda_override : st.lambda);
    for (int e = lane; e < kItemPart; e += 32) {
        if (e >= 36 && cam < 0) break;
        double s = 0.0;
        for (int i = i0; i < i1; ++i) s += P.item_part[kItemPart * (size_t)i + e];
        if (e < 36) {
            if (cam >= 0 && (e % 7) == 0) s += lambda * P.lambda_scale;
            P.S_blk[36 * (size_t)blk + e] = s;
        } else {
            P.bs[6 * (size_t)cam + (e - 36)] = -s;
        }
    }
}

// ------------------------------------------------------------------------------------------
// deterministic sums: chunk partials (fixed tree inside a CTA), then one thread per problem
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_chunk_reduce(DevProblem P, int which)
{
    __shared__ double sh[2][256];
    const int4 ch = P.chunks[blockIdx.x];
    if (!P.lm[ch.x].active && which != 2) return;
    double a = 0.0, b = 0.0;
    const double* src0 = which == 1 ? P.cost_new_pt : P.cost_pt;
    for (int p = ch.y + threadIdx.x; p < ch.z; p += 256) {
        a += src0[p];
        if (which == 1) b += P.scale_pt[p];
    }
    sh[0][threadIdx.x] = a;
    sh[1][threadIdx.x] = b;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            sh[0][threadIdx.x] += sh[0][threadIdx.x + s];
            sh[1][threadIdx.x] += sh[1][threadIdx.x + s];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        P.chunk_part[2 * (size_t)blockIdx.x] = sh[0][0];
        P.chunk_part[2 * (size_t)blockIdx.x + 1] = sh[1][0];
    }
}

// one WARP per problem: the lanes fetch 32 chunk partials at once (one large problem has ~1000 of them: a single thread took one
// round trip after the other), lane 0's additions stay sequential in chunk order
__global__ void __launch_bounds__(128) k_prob_sums(DevProblem P, int which)
{
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (b >= P.n_problems) return;
    if (!P.lm[b].active && which != 2) return;
    double s0 = 0.0, s1 = 0.0;
    const int c1 = P.prob_chunk_ptr[b + 1];
    for (int c = P.prob_chunk_ptr[b]; c < c1; c += 32) {
        const double2 v = c + lane < c1 ? *reinterpret_cast<const double2*>(P.chunk_part + 2 * (size_t)(c + lane)) : make_double2(0.0, 0.0);
        const int n = min(32, c1 - c);
        for (int k = 0; k < n; ++k) {
            s0 += __shfl_sync(0xffffffffu, v.x, k);
            s1 += __shfl_sync(0xffffffffu, v.y, k);
        }
    }
    if (lane) return;
    P.prob_sums[2 * (size_t)b] = s0;
    P.prob_sums[2 * (size_t)b + 1] = s1;
}

// g2o computeLambdaInit + first computeActiveErrors: lambda_0 = tau * max diag(H), chi_0
__global__ void k_lm_init(DevProblem P, double tau, double lambda_init)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= P.n_problems) return;
    LmState& s = P.lm[b];
    double m = __longlong_as_double((long long)P.maxdiag_bits[b]);
    for (int c = P.prob_cam0[b]; c < P.prob_cam0[b + 1]; ++c) {
        const int k = P.diag_blk[c];
        if (k < 0) continue;
        const double* blk = P.S_blk + 36 * (size_t)k;
        for (int r = 0; r < 6; ++r) m = fmax(m, fabs(blk[7 * r]));
    }
    s.chi = P.prob_sums[2 * (size_t)b];
    s.chi_initial = s.chi;
    s.chi_start = s.chi;
    if (!(lambda_init > 0.0)) s.lambda = tau * m;
    s.nu = 2.0;
}

// ------------------------------------------------------------------------------------------
// K3 (small): one CTA per problem.  S (+ rhs as an extra row, so that the forward
// substitution falls out of the factorisation) is assembled densely in shared memory,
// factorised by a right-looking Cholesky, back-substituted by one warp, and the candidate
// poses  T' = exp(d) T  are written straight away.
// ------------------------------------------------------------------------------------------
// The body is shared with the single-window LM kernel (ba_window.cu): problem b, shared memory of small_smem_bytes(n8).
__device__ void solve_small_problem(const DevProblem& P, int b, double* sm, int* s_fail)
{
    LmState& st = P.lm[b];
    const int nf = P.prob_nfree[b];
    const int n = 6 * nf, n8 = (n + 7) & ~7;
    const int ld = small_ld(n8);
    double* const rinv = sm + (size_t)(n8 + 1) * ld;
    const int tid = threadIdx.x, nt = blockDim.x;
    const int cam0 = P.prob_cam0[b], cam1 = P.prob_cam0[b + 1];
    const int tri0 = P.prob_tri0[b];
    const int cur = st.cur;
    if (tid == 0) *s_fail = 0;
    // assemble the lower triangle; rows / columns n .. n8 - 1 are identity padding, row n8 is the right-hand side
    for (int idx = tid; idx < (n8 - n + 1) * ld; idx += nt) sm[n * ld + idx] = 0.0;
    for (int i = n + tid; i < n8; i += nt)
        for (int c = 0; c < n; ++c) sm[c * ld + i] = 0.0; // never read (upper triangle), kept clean
    __syncthreads();
    for (int i = n + tid; i < n8; i += nt) sm[i * ld + i] = 1.0;
    const int nblk = nf * (nf + 1) / 2;
    for (int idx = tid; idx < nblk * 36; idx += nt) {
        const int blk = idx / 36, e = idx - 36 * blk;
        int I = (int)((sqrt(8.0 * (double)blk + 1.0) - 1.0) * 0.5);
        while ((I + 1) * (I + 2) / 2 <= blk) ++I;
        while (I * (I + 1) / 2 > blk) --I;
        const int J = blk - I * (I + 1) / 2;
        const int r = e / 6, c = e - 6 * r;
        sm[(6 * I + r) * ld + 6 * J + c] = P.S_blk[36 * (size_t)(tri0 + blk) + e];
    }
    for (int c = cam0 + tid; c < cam1; c += nt) {
        const int f = P.cam_free[c];
        if (f >= 0)
            for (int r = 0; r < 6; ++r) sm[n8 * ld + 6 * f + r] = P.bs[6 * (size_t)c + r];
    }
    __syncthreads();
    small_chol_forward(sm, n8, ld, rinv, s_fail, tid, nt);
    __syncthreads();
    const int fail = *s_fail;
    if (tid < 32 && !fail) small_backward_warp(sm, n8, ld, rinv, tid);
    __syncthreads();
    // scatter the step, build candidate poses
    const int cand = cur ^ 1;
    for (int c = cam0 + tid; c < cam1; c += nt) {
        const int f = P.cam_free[c];
        double d[6];
        for (int r = 0; r < 6; ++r) d[r] = (f >= 0 && !fail) ? sm[n8 * ld + 6 * f + r] : 0.0;
        for (int r = 0; r < 6; ++r) P.dp[6 * (size_t)c + r] = d[r];
        const double* pc = P.pose[cur] + 7 * (size_t)c;
        double* pn = P.pose[cand] + 7 * (size_t)c;
        if (f >= 0 && !fail) pose_oplus(d, pc, pn);
        else
            for (int r = 0; r < 7; ++r) pn[r] = pc[r];
        write_cam_rt(pn, P.cam_rt[cand] + 12 * (size_t)c);
    }
    if (tid == 0) {
        double ss = 0.0;
        if (!fail)
            for (int i = 0; i < n; ++i) ss += sm[n8 * ld + i] * sm[n8 * ld + i];
        st.dp_sq = ss;
        st.ok2 = fail ? 0 : 1;
    }
}

__global__ void __launch_bounds__(256) k_solve_small(DevProblem P)
{
    extern __shared__ double sm[];
    __shared__ int s_fail;
    const int b = blockIdx.x;
    if (!P.lm[b].active) return;
    solve_small_problem(P, b, sm, &s_fail);
}

// K3 (large, one problem): ba_dense.cu

// ------------------------------------------------------------------------------------------
// K4 update: one thread per point.  Back substitution through the stored records,
//   d_l = L^-T (zb - sum_a G_a^T Jp_a d_p),   P' = P + d_l,
// the point's share of g2o's computeScale  sum_j x_j (lambda x_j + b_j), and the robust
// cost of the candidate state (K5 fused: the new point never leaves registers).
// ------------------------------------------------------------------------------------------
// PHOTO: the candidate cost needs the images: left to k_photo_sample + k_photo_cost (ba_photo.cu)
template <bool PHOTO>
__global__ void __launch_bounds__(128) k_update(DevProblem P)
{
    const int p = P.pt_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.pt_end) return;
    const int prob = P.pt_prob[p];
    const LmState& st = P.lm[prob];
    if (!st.active) return;
    const int cur = st.cur, cand = cur ^ 1;
    const double lambda = st.lambda;
    const double* info = P.ptinfo + kPtInfoDoubles * (size_t)p;
    const double i00 = info[0], i10 = info[1], i11 = info[2], i20 = info[3], i21 = info[4], i22 = info[5];
    const double z0 = info[6], z1 = info[7], z2 = info[8];
    double y0 = z0, y1 = z1, y2 = z2, sc = 0.0;
    const int s0 = P.pt_ptr[p], s1 = P.pt_ptr[p + 1];
    int cam_n = s0 < s1 ? __ldg(P.obs_cam + s0) : 0;
    for (int s = s0; s < s1; ++s) {
        const int cam = cam_n;
        if (s + 1 < s1) cam_n = __ldg(P.obs_cam + s + 1);
        double2 r0, r1, r2, r3, r4, r5;
        {
            const double* r = P.rec + kRecDoubles * (size_t)s;
            ld_global_256(r, r0.x, r0.y, r1.x, r1.y);
            ld_global_256(r + 4, r2.x, r2.y, r3.x, r3.y);
            ld_global_256(r + 8, r4.x, r4.y, r5.x, r5.y);
        }
        const double2* __restrict__ d2 = reinterpret_cast<const double2*>(P.dp + 6 * (size_t)cam); // 48 bytes per pose
        const double2 da = d2[0], db = d2[1], dc = d2[2];
        const double d[6] = { da.x, da.y, db.x, db.y, dc.x, dc.y };
        double ja0[6], ja1[6];
        jtilde(r0.x, r0.y, r1.x, ja0, ja1);
        double t0 = 0.0, t1 = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const double dk = d[k];
            t0 += ja0[k] * dk;
            t1 += ja1[k] * dk;
        }
        // G~ = (r2.x r2.y r3.x ; r3.y r4.x r4.y)
        y0 -= r2.x * t0 + r3.y * t1;
        y1 -= r2.y * t0 + r4.x * t1;
        y2 -= r3.x * t0 + r4.y * t1;
        // pose part of computeScale: d_p . b_p = -sum t~ . (F w r),  F w r = v~ - G~ zb
        const double wr0 = r5.x - (r2.x * z0 + r2.y * z1 + r3.x * z2);
        const double wr1 = r5.y - (r3.y * z0 + r4.x * z1 + r4.y * z2);
        sc -= t0 * wr0 + t1 * wr1;
    }
    const double d0 = i00 * y0 + i10 * y1 + i20 * y2;
    const double d1 = i11 * y1 + i21 * y2;
    const double d2 = i22 * y2;
    sc += y0 * z0 + y1 * z1 + y2 * z2 + lambda * (d0 * d0 + d1 * d1 + d2 * d2);
    const double* pt = P.pts[cur] + 3 * (size_t)p;
    const double X = pt[0] + d0, Y = pt[1] + d1, Z = pt[2] + d2;
    double* pn = P.pts[cand] + 3 * (size_t)p;
    pn[0] = X, pn[1] = Y, pn[2] = Z;
    P.scale_pt[p] = sc;
    if (PHOTO) return;
    const double* __restrict__ cam_rt = P.cam_rt[cand];
    double cost = 0.0;
    cam_n = s0 < s1 ? __ldg(P.obs_cam + s0) : 0;
    float2 uv_n = s0 < s1 ? __ldg(P.obs_uv + s0) : make_float2(0.f, 0.f);
    for (int s = s0; s < s1; ++s) {
        const int cam = cam_n;
        const float2 uv = uv_n;
        if (s + 1 < s1) cam_n = __ldg(P.obs_cam + s + 1), uv_n = __ldg(P.obs_uv + s + 1);
        double rt[12];
        load_rt(cam_rt, cam, rt);
        const double x = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
        const double y = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
        const double z = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
        const double iz = 1.0 / z;
        const double ru = P.fx * (x * iz) + P.cx - (double)uv.x;
        const double rv = P.fy * (y * iz) + P.cy - (double)uv.y;
        double w;
        cost += huber_rho(ru * ru + rv * rv, P.huber_delta, w);
    }
    P.cost_new_pt[p] = cost;
}

// ------------------------------------------------------------------------------------------
// LM decision, one thread per problem: g2o OptimizationAlgorithmLevenberg::solve, the part
// after the candidate has been evaluated.  Accepting flips `cur`; rejecting leaves the state.
// ------------------------------------------------------------------------------------------
__global__ void k_decide(DevProblem P)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= P.n_problems) return;
    lm_decide(P, b);
}

// ------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------
static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

int launch_state_init(const DevProblem& P, int iterations, int max_trials, double lambda_init, cudaStream_t st)
{
    k_state_init<<<cdiv(P.n_problems, 128), 128, 0, st>>>(P, iterations, max_trials, lambda_init);
    return 1;
}

int launch_cam_rt(const DevProblem& P, int which, cudaStream_t st)
{
    k_cam_rt<<<cdiv(P.n_cams, 128), 128, 0, st>>>(P, which);
    return 1;
}

int launch_gather_current(const DevProblem& P, double* poses_out, double* pts_out, cudaStream_t st)
{
    const long long n = 7ll * P.n_cams + 3ll * P.n_points;
    if (n <= 0) return 0;
    k_gather_current<<<cdiv(n, 256), 256, 0, st>>>(P, poses_out, pts_out);
    return 1;
}

int launch_linearize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st)
{
    const int np = P.pt_end - P.pt_begin;
    if (np <= 0) return 0;
    k_linearize<<<cdiv(np, 128), 128, 0, st>>>(P, init_pass, lambda_override);
    return 1;
}

int launch_eval_cost(const DevProblem& P, cudaStream_t st)
{
    const int np = P.pt_end - P.pt_begin;
    if (np <= 0) return 0;
    k_eval_cost<<<cdiv(np, 128), 128, 0, st>>>(P);
    return 1;
}

int launch_schur_blocks(const DevProblem& P, int diag_only, cudaStream_t st)
{
    const long long warps = diag_only ? P.n_diag_items : P.n_items;
    if (warps <= 0) return 0;
    static PerDeviceOnce once;
    if (!once.run([] {
            return cudaFuncSetAttribute(k_schur_blocks<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSbBytes) == cudaSuccess
                && cudaFuncSetAttribute(k_schur_blocks<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSbBytes) == cudaSuccess;
        }))
        return 0;
    if (P.ph_sum) k_schur_blocks<true><<<cdiv(warps * 32, kSbThreads), kSbThreads, kSbBytes, st>>>(P, diag_only);
    else k_schur_blocks<false><<<cdiv(warps * 32, kSbThreads), kSbThreads, kSbBytes, st>>>(P, diag_only);
    return 1;
}

int launch_block_finalize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st)
{
    if (P.n_blocks <= 0) return 0;
    k_block_finalize<<<cdiv((long long)P.n_blocks * 32, 128), 128, 0, st>>>(P, init_pass, lambda_override);
    return 1;
}

int launch_reduce(const DevProblem& P, int which, cudaStream_t st)
{
    int n = 0;
    if (P.n_chunks > 0) {
        k_chunk_reduce<<<P.n_chunks, 256, 0, st>>>(P, which);
        ++n;
    }
    k_prob_sums<<<cdiv(P.n_problems, 4), 128, 0, st>>>(P, which);
    return n + 1;
}

int launch_lm_init(const DevProblem& P, double tau, double lambda_init, cudaStream_t st)
{
    k_lm_init<<<cdiv(P.n_problems, 128), 128, 0, st>>>(P, tau, lambda_init);
    return 1;
}

static int win_sm_count_hint()
{
    static PerDeviceOnce once;
    return once.sm_count();
}

int launch_solve_small(const DevProblem& P, int max_n, cudaStream_t st)
{
    const int n8 = (max_n + 7) & ~7;
    const size_t smem = small_smem_bytes(n8);
    static PerDeviceOnce once; // the largest system the kernel takes, once per device
    constexpr size_t kMaxSmem = small_smem_bytes(kSmallN8);
    if (smem > kMaxSmem) return 0;
    if (!once.run([] { return cudaFuncSetAttribute(k_solve_small, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmem) == cudaSuccess; })) return 0;
    // a batch of small systems: 64 threads per CTA, so that all of 1024 problems are resident at once (8 CTAs per SM instead of
    // 4 at 104 registers: one wave instead of two); a few systems: more threads per system.  Same arithmetic either way.
    const int threads = max_n <= 64 ? (P.n_problems >= 4 * win_sm_count_hint() ? 64 : 128) : 256;
    k_solve_small<<<P.n_problems, threads, smem, st>>>(P);
    return 1;
}

int launch_update(const DevProblem& P, cudaStream_t st)
{
    const int np = P.pt_end - P.pt_begin;
    if (np <= 0) return 0;
    if (P.ph_sum) k_update<true><<<cdiv(np, 128), 128, 0, st>>>(P);
    else k_update<false><<<cdiv(np, 128), 128, 0, st>>>(P);
    return 1;
}

int launch_decide(const DevProblem& P, cudaStream_t st)
{
    k_decide<<<cdiv(P.n_problems, 128), 128, 0, st>>>(P);
    return 1;
}

} // namespace visoba
