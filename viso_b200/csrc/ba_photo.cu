// ba_photo.cu — the reference adjuster's own edge, EdgeDirectProjection (reference include/bundle_adjuster.h:58-106), as
// a residual type of the Levenberg-Marquardt loop.
//
//   residual (16)   r_k = I_src(u_s + o_k) - I_tgt(u_t + o_k),  o_k the 4 x 4 patch i, j = -2..1   (:83-97)
//                   u_t = K (R p + t) / z with the pose VERTEX (:73-78), u_s = srcFrame->Project(p) with the keyframe's own
//                   constant pose (:79, keyframe.h:82-89), sample coordinates cast to float (:86-89), information I_16
//   disabled edge   a sample outside either image: setLevel(1) (:91-95) -> the edge contributes nothing at that state
//   Jacobian        the reference defines none (g2o would difference numerically); here analytic:
//                     d r_k / d xi = -g_t,k^T B          B = dPixeldXi = F J~      (viso.cpp:640-658)
//                     d r_k / d p  =  g_s,k^T A_s - g_t,k^T A_t      A = first three columns of dPixeldXi times R
//                   with g the image gradient of keyframe.h:59-64 (central differences of the bilinear sampler).
//
// How it folds into the existing Schur machinery.  With the 2 x 2 sums over the 16 samples
//     C_ss = sum g_s g_s^T,  C_st = sum g_s g_t^T,  C_tt = sum g_t g_t^T,  c_s = sum g_s r,  c_t = sum g_t r
// the edge's normal-equation blocks are
//     H_pp = J~^T (F C_tt F) J~          H_pl = J~^T W^,  W^ = F (C_tt A_t - C_st^T A_s)          b_p = J~^T F c_t
//     H_ll = A_s^T C_ss A_s - A_s^T C_st A_t - A_t^T C_st^T A_s + A_t^T C_tt A_t                   b_l = A_t^T c_t - A_s^T c_s
// i.e. exactly the shape of the reprojection edge (ba_kernels.cu) with the scalar weight w F^2 of the diagonal block
// replaced by the symmetric 2 x 2 matrix F C_tt F.  So the per-observation record {x/z, y/z, 1/z, -, G~ = W^ L^-T, v~} is
// kept, the 2 x 2 middle goes to a side array (ph_mid), and k_schur_blocks / k_update run unchanged but for that term and
// the candidate cost (template flag).  Three kernels are new:
//     k_photo_sample     16 lanes per edge: the samples, gradients and their sums (fixed butterfly: deterministic)
//     k_photo_linearize  thread per point: H_ll, b_l, 3 x 3 Cholesky, records            (the photometric k_linearize)
//     k_photo_cost       thread per point: robustified cost of its edges from the sums
#include "ba_math.cuh"
#include "image_math.cuh"

namespace visoba {

using visoimg::DevImage;

namespace {

constexpr int kPhSum = 16; // doubles per edge: C_ss (3) C_st (4) C_tt (3) c_s (2) c_t (2) chi2 valid

__device__ __forceinline__ double sum16(double v)
{ // total over the 16 lanes of the edge, the same fixed tree in every lane
#pragma unroll
    for (int off = 8; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

__device__ __forceinline__ int point_of_obs(const DevProblem& P, int s)
{ // the CSR row that holds observation s (pt_ptr is non-decreasing): last p with pt_ptr[p] <= s
    int lo = 0, hi = P.n_points - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(P.pt_ptr + mid) <= s) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ int src_of(float2 uv) { return __float_as_int(uv.x); } // the source keyframe rides in the measurement slot

} // namespace

// which: 0 = every problem's current state, 1 = its candidate state.  full = 0: residual only (cost evaluation).
__global__ void __launch_bounds__(128) k_photo_sample(DevProblem P, int which, int full, int even_if_inactive)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int s = t >> 4, k = t & 15;
    if (s >= P.n_obs) return; // whole 16-lane groups leave together
    const int p = point_of_obs(P, s);
    const LmState& st = P.lm[P.pt_prob[p]];
    if (!st.active && !even_if_inactive) return;
    const int buf = which ? st.cur ^ 1 : st.cur;
    const int dst = __ldg(P.obs_cam + s), src = src_of(__ldg(P.obs_uv + s));
    const DevImage* frames = reinterpret_cast<const DevImage*>(P.ph_frames);
    const DevImage tgt = frames[dst], sim = frames[src];
    const double* __restrict__ pt = P.pts[buf] + 3 * (size_t)p;
    const double X = pt[0], Y = pt[1], Z = pt[2];
    const double* __restrict__ rt = P.cam_rt[buf] + 12 * (size_t)dst; // T->estimate(), :73-75
    const double x = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
    const double y = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
    const double z = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
    const double tu = (P.fx * x + 0.0 * y + P.cx * z) / z, tv = (0.0 * x + P.fy * y + P.cy * z) / z; // :77-78
    const double* __restrict__ rs = P.ph_src_rt + 12 * (size_t)src; // srcFrame->Project(p), :79
    const double xs = rs[0] * X + rs[1] * Y + rs[2] * Z + rs[9];
    const double ys = rs[3] * X + rs[4] * Y + rs[5] * Z + rs[10];
    const double zs = rs[6] * X + rs[7] * Y + rs[8] * Z + rs[11];
    const double su = 1.0 * ((xs / zs) * P.fx + P.cx), sv = 1.0 * ((ys / zs) * P.fy + P.cy);
    const int i = (k & 3) - 2, j = (k >> 2) - 2;                                // :83-84
    const float u1 = (float)(tu + (double)i), v1 = (float)(tv + (double)j);     // :86-87
    const float u2 = (float)(su + (double)i), v2 = (float)(sv + (double)j);     // :88-89
    const bool outside = !visoimg::is_inside(tgt, u1, v1) || !visoimg::is_inside(sim, u2, v2);
    const unsigned half = (threadIdx.x & 16) ? 0xffff0000u : 0x0000ffffu;
    const bool valid = (__ballot_sync(0xffffffffu, outside) & half) == 0u;      // :91-95
    double* out = P.ph_sum + kPhSum * (size_t)s;
    if (!valid) {
        out[k] = 0.0;
        return;
    }
    const double r = visoimg::pixel_value(sim, u2, v2) - visoimg::pixel_value(tgt, u1, v1); // :97
    double v[kPhSum];
#pragma unroll
    for (int q = 0; q < kPhSum; ++q) v[q] = 0.0;
    if (full) {
        double gsx, gsy, gtx, gty;
        visoimg::gradient(sim, u2, v2, gsx, gsy);
        visoimg::gradient(tgt, u1, v1, gtx, gty);
        v[0] = gsx * gsx, v[1] = gsx * gsy, v[2] = gsy * gsy;                               // C_ss: xx xy yy
        v[3] = gsx * gtx, v[4] = gsx * gty, v[5] = gsy * gtx, v[6] = gsy * gty;             // C_st: row s, column t
        v[7] = gtx * gtx, v[8] = gtx * gty, v[9] = gty * gty;                               // C_tt
        v[10] = gsx * r, v[11] = gsy * r, v[12] = gtx * r, v[13] = gty * r;                 // c_s, c_t
    }
    v[14] = r * r;
    double mine = 0.0;
#pragma unroll
    for (int q = 0; q < 15; ++q) {
        const double tot = sum16(v[q]);
        if (k == q) mine = tot;
    }
    out[k] = k == 15 ? 1.0 : mine;
}

// The photometric k_linearize: one thread per point (see ba_kernels.cu::k_linearize for the shape of the passes).
__global__ void __launch_bounds__(128) k_photo_linearize(DevProblem P, int init_pass, double lambda_override)
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
    const float2 no_uv = make_float2(0.f, 0.f);

    double h00 = 0, h10 = 0, h11 = 0, h20 = 0, h21 = 0, h22 = 0, b0 = 0, b1 = 0, b2 = 0, cost = 0;
    for (int s = s0; s < s1; ++s) {
        const double* __restrict__ q = P.ph_sum + kPhSum * (size_t)s;
        if (q[15] == 0.0) continue; // the edge disabled itself at this state
        double w;
        cost += huber_rho(q[14], P.huber_delta, w);
        ObsGeom gt, gs; // only the Jacobian rows jl0 / jl1 (= A with fx, fy) are used
        obs_geometry(P, cam_rt + 12 * (size_t)__ldg(P.obs_cam + s), X, Y, Z, no_uv, gt);
        obs_geometry(P, P.ph_src_rt + 12 * (size_t)src_of(__ldg(P.obs_uv + s)), X, Y, Z, no_uv, gs);
        // rows of D = C_ss A_s - C_st A_t and E = C_tt A_t - C_st^T A_s  (2 x 3 each):  H_ll = A_s^T D + A_t^T E
        double D0[3], D1[3], E0[3], E1[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            D0[c] = q[0] * gs.jl0[c] + q[1] * gs.jl1[c] - (q[3] * gt.jl0[c] + q[4] * gt.jl1[c]);
            D1[c] = q[1] * gs.jl0[c] + q[2] * gs.jl1[c] - (q[5] * gt.jl0[c] + q[6] * gt.jl1[c]);
            E0[c] = q[7] * gt.jl0[c] + q[8] * gt.jl1[c] - (q[3] * gs.jl0[c] + q[5] * gs.jl1[c]);
            E1[c] = q[8] * gt.jl0[c] + q[9] * gt.jl1[c] - (q[4] * gs.jl0[c] + q[6] * gs.jl1[c]);
        }
#define HLL(a, c) (gs.jl0[a] * D0[c] + gs.jl1[a] * D1[c] + gt.jl0[a] * E0[c] + gt.jl1[a] * E1[c])
        h00 += w * HLL(0, 0), h10 += w * HLL(1, 0), h11 += w * HLL(1, 1);
        h20 += w * HLL(2, 0), h21 += w * HLL(2, 1), h22 += w * HLL(2, 2);
#undef HLL
        // b_l = -J_l^T r = A_t^T c_t - A_s^T c_s
        b0 += w * (gt.jl0[0] * q[12] + gt.jl1[0] * q[13] - (gs.jl0[0] * q[10] + gs.jl1[0] * q[11]));
        b1 += w * (gt.jl0[1] * q[12] + gt.jl1[1] * q[13] - (gs.jl0[1] * q[10] + gs.jl1[1] * q[11]));
        b2 += w * (gt.jl0[2] * q[12] + gt.jl1[2] * q[13] - (gs.jl0[2] * q[10] + gs.jl1[2] * q[11]));
    }
    P.cost_pt[p] = cost;
    double i00 = 0, i10 = 0, i11 = 0, i20 = 0, i21 = 0, i22 = 0, z0 = 0, z1 = 0, z2 = 0;
    if (init_pass) { // g2o computeLambdaInit: the point part of max diag(H); max is order-free
        const double m = fmax(h00, fmax(h11, h22));
        if (m > 0.0) atomicMax(P.maxdiag_bits + prob, (unsigned long long)__double_as_longlong(m));
    } else {
        h00 += lambda, h11 += lambda, h22 += lambda;
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

    for (int s = s0; s < s1; ++s) {
        const double* __restrict__ q = P.ph_sum + kPhSum * (size_t)s;
        ObsGeom gt, gs;
        obs_geometry(P, cam_rt + 12 * (size_t)__ldg(P.obs_cam + s), X, Y, Z, no_uv, gt);
        obs_geometry(P, P.ph_src_rt + 12 * (size_t)src_of(__ldg(P.obs_uv + s)), X, Y, Z, no_uv, gs);
        double w = 0.0;
        if (q[15] != 0.0) huber_rho(q[14], P.huber_delta, w);
        // W^ / F = w (C_tt A_t - C_st^T A_s)   (2 x 3),   G = W^ L^-T
        double E0[3], E1[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            E0[c] = w * (q[7] * gt.jl0[c] + q[8] * gt.jl1[c] - (q[3] * gs.jl0[c] + q[5] * gs.jl1[c]));
            E1[c] = w * (q[8] * gt.jl0[c] + q[9] * gt.jl1[c] - (q[4] * gs.jl0[c] + q[6] * gs.jl1[c]));
        }
        const double g00 = E0[0] * i00;
        const double g01 = E0[0] * i10 + E0[1] * i11;
        const double g02 = E0[0] * i20 + E0[1] * i21 + E0[2] * i22;
        const double g10 = E1[0] * i00;
        const double g11 = E1[0] * i10 + E1[1] * i11;
        const double g12 = E1[0] * i20 + E1[1] * i21 + E1[2] * i22;
        // v~ / F = -w c_t + G zb:   bs_I = -sum J~^T v~ = b_p - H_pl H_ll^-1 b_l
        const double v0 = -w * q[12] + g00 * z0 + g01 * z1 + g02 * z2;
        const double v1 = -w * q[13] + g10 * z0 + g11 * z1 + g12 * z2;
        double2* r = reinterpret_cast<double2*>(P.rec + kRecDoubles * (size_t)s);
        r[0] = make_double2(gt.xn, gt.yn);
        r[1] = make_double2(gt.iz, 0.0);
        r[2] = make_double2(P.fx * g00, P.fx * g01);
        r[3] = make_double2(P.fx * g02, P.fy * g10);
        r[4] = make_double2(P.fy * g11, P.fy * g12);
        r[5] = make_double2(P.fx * v0, P.fy * v1);
        double* m = P.ph_mid + 4 * (size_t)s; // the 2 x 2 middle of the diagonal block:  w F C_tt F
        m[0] = w * P.fx * P.fx * q[7], m[1] = w * P.fx * P.fy * q[8], m[2] = w * P.fy * P.fy * q[9], m[3] = 0.0;
    }
}

// robustified cost of a point's edges from the sums k_photo_sample left (which: 0 -> cost_pt, 1 -> cost_new_pt)
__global__ void __launch_bounds__(128) k_photo_cost(DevProblem P, int which, int even_if_inactive)
{
    const int p = P.pt_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.pt_end) return;
    if (!P.lm[P.pt_prob[p]].active && !even_if_inactive) return;
    double cost = 0.0;
    for (int s = P.pt_ptr[p]; s < P.pt_ptr[p + 1]; ++s) {
        const double* __restrict__ q = P.ph_sum + kPhSum * (size_t)s;
        double w;
        if (q[15] != 0.0) cost += huber_rho(q[14], P.huber_delta, w);
    }
    (which ? P.cost_new_pt : P.cost_pt)[p] = cost;
}

int launch_photo_sample(const DevProblem& P, int which, int full, int even_if_inactive, cudaStream_t st)
{
    if (P.n_obs <= 0) return 0;
    k_photo_sample<<<(unsigned)((P.n_obs * 16 + 127) / 128), 128, 0, st>>>(P, which, full, even_if_inactive);
    return 1;
}

int launch_photo_linearize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st)
{
    const int n = P.pt_end - P.pt_begin;
    if (n <= 0) return 0;
    k_photo_linearize<<<(n + 127) / 128, 128, 0, st>>>(P, init_pass, lambda_override);
    return 1;
}

int launch_photo_cost(const DevProblem& P, int which, int even_if_inactive, cudaStream_t st)
{
    const int n = P.pt_end - P.pt_begin;
    if (n <= 0) return 0;
    k_photo_cost<<<(n + 127) / 128, 128, 0, st>>>(P, which, even_if_inactive);
    return 1;
}

size_t photo_sum_doubles() { return kPhSum; }

} // namespace visoba
