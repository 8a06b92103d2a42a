// ba_math.cuh — device-side math shared by the general kernels (ba_kernels.cu) and the window
// fast path (ba_window.cu).  Reference semantics (citations relative to /root/reference):
//   projection          include/keyframe.h:82-89          Pc = R Pw + t, u = fx x/z + cx
//   2x6 pose Jacobian   src/viso.cpp:640-658 (dPixeldXi)  columns (v, w), left perturbation
//   2x3 point Jacobian  first three columns of dPixeldXi times R (absent in the reference)
//   pose update         include/bundle_adjuster.h:51-55   T <- exp(d) T
#ifndef VISO_BA_MATH_CUH
#define VISO_BA_MATH_CUH

#include "ba_device.h"

#include <float.h>
#include <math.h>

namespace visoba {

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double huber_rho(double s, double delta, double& w)
{ // g2o RobustKernelHuber::robustify (rho[0], rho[1]); delta <= 0 -> plain L2
    w = 1.0;
    if (delta <= 0.0) return s;
    const double dsqr = delta * delta;
    if (s <= dsqr) return s;
    const double sq = sqrt(s);
    w = delta / sq;
    return 2.0 * sq * delta - dsqr;
}

__device__ __forceinline__ void quat_to_R(const double* q, double* R)
{
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    R[0] = 1.0 - 2.0 * (y * y + z * z);
    R[1] = 2.0 * (x * y - z * w);
    R[2] = 2.0 * (x * z + y * w);
    R[3] = 2.0 * (x * y + z * w);
    R[4] = 1.0 - 2.0 * (x * x + z * z);
    R[5] = 2.0 * (y * z - x * w);
    R[6] = 2.0 * (x * z - y * w);
    R[7] = 2.0 * (y * z + x * w);
    R[8] = 1.0 - 2.0 * (x * x + y * y);
}

// Sophus SE3d::exp(xi) * T, xi = (v, w); writes the new pose (q xyzw, t)
__device__ __forceinline__ void pose_oplus(const double* xi, const double* pose, double* out)
{
    const double wx = xi[3], wy = xi[4], wz = xi[5];
    const double th2 = wx * wx + wy * wy + wz * wz;
    double im, re, a, b;
    if (th2 < 1e-10) {
        const double th4 = th2 * th2;
        im = 0.5 - th2 / 48.0 + th4 / 3840.0;
        re = 1.0 - th2 / 8.0 + th4 / 384.0;
        a = 0.5 - th2 / 24.0 + th4 / 720.0;
        b = 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0;
    } else {
        const double th = sqrt(th2);
        double sh, ch, s1, c1;
        sincos(0.5 * th, &sh, &ch);
        sincos(th, &s1, &c1);
        im = sh / th;
        re = ch;
        a = (1.0 - c1) / th2;
        b = (th - s1) / (th2 * th);
    }
    const double dq[4] = { im * wx, im * wy, im * wz, re };
    const double vx = xi[0], vy = xi[1], vz = xi[2];
    const double c1x = wy * vz - wz * vy, c1y = wz * vx - wx * vz, c1z = wx * vy - wy * vx;
    const double c2x = wy * c1z - wz * c1y, c2y = wz * c1x - wx * c1z, c2z = wx * c1y - wy * c1x;
    const double dt0 = vx + a * c1x + b * c2x;
    const double dt1 = vy + a * c1y + b * c2y;
    const double dt2 = vz + a * c1z + b * c2z;
    double dR[9];
    quat_to_R(dq, dR);
    const double bx = pose[0], by = pose[1], bz = pose[2], bw = pose[3];
    const double qx = dq[3] * bx + dq[0] * bw + dq[1] * bz - dq[2] * by;
    const double qy = dq[3] * by - dq[0] * bz + dq[1] * bw + dq[2] * bx;
    const double qz = dq[3] * bz + dq[0] * by - dq[1] * bx + dq[2] * bw;
    const double qw = dq[3] * bw - dq[0] * bx - dq[1] * by - dq[2] * bz;
    const double n = 1.0 / sqrt(qx * qx + qy * qy + qz * qz + qw * qw);
    out[0] = qx * n;
    out[1] = qy * n;
    out[2] = qz * n;
    out[3] = qw * n;
    const double tx = pose[4], ty = pose[5], tz = pose[6];
    out[4] = dR[0] * tx + dR[1] * ty + dR[2] * tz + dt0;
    out[5] = dR[3] * tx + dR[4] * ty + dR[5] * tz + dt1;
    out[6] = dR[6] * tx + dR[7] * ty + dR[8] * tz + dt2;
}

__device__ __forceinline__ void write_cam_rt(const double* pose, double* rt)
{
    double R[9];
    quat_to_R(pose, R);
#pragma unroll
    for (int i = 0; i < 9; ++i) rt[i] = R[i];
    rt[9] = pose[4];
    rt[10] = pose[5];
    rt[11] = pose[6];
}

struct ObsGeom {
    double xn, yn, iz, ru, rv, w, rho;
    double jl0[3], jl1[3]; // rows of the 2x3 point Jacobian (with fx, fy)
};

__device__ __forceinline__ void obs_geometry(const DevProblem& P, const double* __restrict__ rt,
    double X, double Y, double Z, float2 uv, ObsGeom& g)
{
    const double x = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
    const double y = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
    const double z = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
    g.iz = 1.0 / z;
    g.xn = x * g.iz;
    g.yn = y * g.iz;
    g.ru = P.fx * g.xn + P.cx - (double)uv.x;
    g.rv = P.fy * g.yn + P.cy - (double)uv.y;
    g.rho = huber_rho(g.ru * g.ru + g.rv * g.rv, P.huber_delta, g.w);
    const double a0 = P.fx * g.iz, a2 = -P.fx * g.xn * g.iz;
    const double b1 = P.fy * g.iz, b2 = -P.fy * g.yn * g.iz;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        g.jl0[k] = a0 * rt[k] + a2 * rt[6 + k];
        g.jl1[k] = b1 * rt[3 + k] + b2 * rt[6 + k];
    }
}

__device__ __forceinline__ void jtilde(double xn, double yn, double iz, double (&r0)[6], double (&r1)[6])
{ // viso.cpp:653-655 without the fx / fy factors
    r0[0] = iz, r0[1] = 0.0, r0[2] = -xn * iz, r0[3] = -xn * yn, r0[4] = 1.0 + xn * xn, r0[5] = -yn;
    r1[0] = 0.0, r1[1] = iz, r1[2] = -yn * iz, r1[3] = -(1.0 + yn * yn), r1[4] = xn * yn, r1[5] = xn;
}

// LM decision of problem b: g2o OptimizationAlgorithmLevenberg::solve, the part after the candidate has been evaluated
// (prob_sums[b] = {candidate cost, point part of computeScale}).  Accepting flips `cur`; rejecting leaves the state.
__device__ __forceinline__ void lm_decide(const DevProblem& P, int b)
{
    LmState& s = P.lm[b];
    if (!s.active) return;
    if (s.q == 0) {
        s.chi_start = s.chi;
        s.accepted = 0;
    }
    double chi_new = P.prob_sums[2 * (size_t)b];
    if (!s.ok2) chi_new = DBL_MAX;
    const double scale = P.prob_sums[2 * (size_t)b + 1] + s.lambda * s.dp_sq + 1e-3;
    const double rho = (s.chi - chi_new) / scale;
    s.rho = rho;
    bool lambda_bad = false;
    if (rho > 0.0 && isfinite(chi_new)) {
        double alpha = 1.0 - (2.0 * rho - 1.0) * (2.0 * rho - 1.0) * (2.0 * rho - 1.0);
        alpha = fmin(alpha, 2.0 / 3.0);
        s.lambda *= fmax(1.0 / 3.0, alpha);
        s.nu = 2.0;
        s.chi = chi_new;
        s.cur ^= 1;
        s.accepted = 1;
    } else {
        s.lambda *= s.nu;
        s.nu *= 2.0;
        if (!isfinite(s.lambda)) lambda_bad = true;
    }
    s.q += 1;
    s.trials += 1;
    if (!lambda_bad && rho < 0.0 && s.q < s.max_trials) return; // another trial, same iteration
    if (P.iter_stats && b == 0) {
        IterStat& r = P.iter_stats[s.it];
        r.cost_before = s.chi_start;
        r.cost_after = s.chi;
        r.lambda = s.lambda;
        r.trials = s.q;
        r.accepted = s.accepted;
    }
    const bool term = (s.q == s.max_trials) || rho == 0.0 || lambda_bad;
    s.it += 1;
    s.q = 0;
    if (term) s.terminated = 1;
    if (term || s.it >= s.target_it) {
        s.active = 0;
        atomicSub(P.n_active, 1);
    }
}

} // namespace visoba
#endif
