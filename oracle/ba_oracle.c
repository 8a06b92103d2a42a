/*
 * ba_oracle.c — plain-C fp64 CPU oracle for the viso bundle-adjustment path.
 *
 * TEST INFRASTRUCTURE ONLY (see ba_oracle.h).  PARITY UNPINNED: the reference
 * adjuster does not compile and ships no tests; this restates its conventions
 * and g2o's published LM / Schur / Huber algorithm.  Citations are relative to
 * /root/reference.
 *
 * Conventions followed:
 *   pose direction + projection   include/keyframe.h:82-89
 *   shared intrinsics             src/main.cpp:14-17, src/viso.cpp:10
 *   2x6 pose Jacobian, (v,w) order src/viso.cpp:640-658
 *   pose update  T <- exp(d) * T  include/bundle_adjuster.h:51-55, src/viso.cpp:737
 *   point update P <- P + d       include/bundle_adjuster.h:34-38
 *   edge = (point, pose)          include/bundle_adjuster.h:72-73, :158-159
 *   6/3 blocks, points marginalised include/bundle_adjuster.h:108, :128
 *   dense solve of the pose system  include/bundle_adjuster.h:113-114
 *   Levenberg-Marquardt             include/bundle_adjuster.h:116-117 (g2o policy)
 */
#include "ba_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

void oracle_default_cfg(oracle_cfg* c)
{
    c->fx = 517.3; /* src/main.cpp:14-17 */
    c->fy = 516.5;
    c->cx = 325.1;
    c->cy = 249.7;
    c->huber_delta = 0.0;
    c->tau = 1e-5;
    c->lambda_init = 0.0;
    c->max_trials = 10;
    c->threads = 1;
}

void oracle_quat_to_R(const double q[4], double R[9])
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

static void transform(const double R[9], const double t[3], const double P[3], double Pc[3])
{
    Pc[0] = R[0] * P[0] + R[1] * P[1] + R[2] * P[2] + t[0];
    Pc[1] = R[3] * P[0] + R[4] * P[1] + R[5] * P[2] + t[1];
    Pc[2] = R[6] * P[0] + R[7] * P[1] + R[8] * P[2] + t[2];
}

/* keyframe.h:82-89 */
void oracle_project(const oracle_cfg* c, const double pose[7], const double Pw[3], double uv[2],
                    double Pc_out[3])
{
    double R[9], Pc[3];
    oracle_quat_to_R(pose, R);
    transform(R, pose + 4, Pw, Pc);
    uv[0] = c->fx * (Pc[0] / Pc[2]) + c->cx;
    uv[1] = c->fy * (Pc[1] / Pc[2]) + c->cy;
    if (Pc_out) {
        Pc_out[0] = Pc[0];
        Pc_out[1] = Pc[1];
        Pc_out[2] = Pc[2];
    }
}

/* viso.cpp:640-658 (entries at :653-655), scale = 1 */
void oracle_jac_pose(const oracle_cfg* c, const double Pc[3], double J[12])
{
    const double x = Pc[0], y = Pc[1], z = Pc[2];
    const double fx = c->fx, fy = c->fy;
    const double zz = z * z, xy = x * y;
    J[0] = fx / z;
    J[1] = 0.0;
    J[2] = -fx * x / zz;
    J[3] = -fx * xy / zz;
    J[4] = fx + fx * x * x / zz;
    J[5] = -fx * y / z;
    J[6] = 0.0;
    J[7] = fy / z;
    J[8] = -fy * y / zz;
    J[9] = -fy - fy * y * y / zz;
    J[10] = fy * xy / zz;
    J[11] = fy * x / z;
}

/* d(u,v)/dPw = J_proj(2x3) * R, J_proj = first three columns of dPixeldXi */
void oracle_jac_point(const oracle_cfg* c, const double Pc[3], const double R[9], double J[6])
{
    const double x = Pc[0], y = Pc[1], z = Pc[2];
    const double zz = z * z;
    const double a0 = c->fx / z, a2 = -c->fx * x / zz;
    const double b1 = c->fy / z, b2 = -c->fy * y / zz;
    for (int k = 0; k < 3; ++k) {
        J[k] = a0 * R[k] + a2 * R[6 + k];
        J[3 + k] = b1 * R[3 + k] + b2 * R[6 + k];
    }
}

/* Sophus SE3d::exp semantics (restated): q = Exp(w), t = V(w) * v */
void oracle_se3_exp(const double xi[6], double pose[7])
{
    const double wx = xi[3], wy = xi[4], wz = xi[5];
    const double th2 = wx * wx + wy * wy + wz * wz;
    const double th = sqrt(th2);
    double imag, real, a, b;
    if (th2 < 1e-10) {
        const double th4 = th2 * th2;
        imag = 0.5 - th2 / 48.0 + th4 / 3840.0;
        real = 1.0 - th2 / 8.0 + th4 / 384.0;
        a = 0.5 - th2 / 24.0 + th4 / 720.0;
        b = 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0;
    } else {
        const double h = 0.5 * th;
        imag = sin(h) / th;
        real = cos(h);
        a = (1.0 - cos(th)) / th2;
        b = (th - sin(th)) / (th2 * th);
    }
    pose[0] = imag * wx;
    pose[1] = imag * wy;
    pose[2] = imag * wz;
    pose[3] = real;
    /* V v = v + a (w x v) + b (w x (w x v)) */
    const double vx = xi[0], vy = xi[1], vz = xi[2];
    const double c1x = wy * vz - wz * vy, c1y = wz * vx - wx * vz, c1z = wx * vy - wy * vx;
    const double c2x = wy * c1z - wz * c1y, c2y = wz * c1x - wx * c1z, c2z = wx * c1y - wy * c1x;
    pose[4] = vx + a * c1x + b * c2x;
    pose[5] = vy + a * c1y + b * c2y;
    pose[6] = vz + a * c1z + b * c2z;
}

/* bundle_adjuster.h:51-55: estimate <- SE3::exp(update) * estimate */
void oracle_pose_oplus(const double xi[6], double pose[7])
{
    double d[7], R[9];
    oracle_se3_exp(xi, d);
    oracle_quat_to_R(d, R);
    /* q <- dq * q (Hamilton product, xyzw) */
    const double ax = d[0], ay = d[1], az = d[2], aw = d[3];
    const double bx = pose[0], by = pose[1], bz = pose[2], bw = pose[3];
    double qx = aw * bx + ax * bw + ay * bz - az * by;
    double qy = aw * by - ax * bz + ay * bw + az * bx;
    double qz = aw * bz + ax * by - ay * bx + az * bw;
    double qw = aw * bw - ax * bx - ay * by - az * bz;
    const double n = 1.0 / sqrt(qx * qx + qy * qy + qz * qz + qw * qw);
    pose[0] = qx * n;
    pose[1] = qy * n;
    pose[2] = qz * n;
    pose[3] = qw * n;
    /* t <- dR * t + dt */
    const double tx = pose[4], ty = pose[5], tz = pose[6];
    pose[4] = R[0] * tx + R[1] * ty + R[2] * tz + d[4];
    pose[5] = R[3] * tx + R[4] * ty + R[5] * tz + d[5];
    pose[6] = R[6] * tx + R[7] * ty + R[8] * tz + d[6];
}

/* g2o RobustKernelHuber::robustify: rho[0], rho[1] */
double oracle_huber(double s, double delta, double* w)
{
    if (delta <= 0.0) {
        *w = 1.0;
        return s;
    }
    const double dsqr = delta * delta;
    if (s <= dsqr) {
        *w = 1.0;
        return s;
    }
    const double sq = sqrt(s);
    *w = delta / sq;
    return 2.0 * sq * delta - dsqr;
}

/* ------------------------------------------------------------------------ */

typedef struct {
    int32_t n_poses, n_points, n_free, n;
    int64_t n_obs;
    int32_t* free_index; /* pose -> free slot or -1 */
    int64_t* pt_ptr;     /* CSR by point, n_points + 1 */
    int64_t* perm;       /* sorted position -> original observation */
    double* Rt;          /* n_poses x 12: R row-major then t */
    double* W;           /* n_obs x 18 (6x3 row-major), sorted order */
    double* Hpp;         /* n_poses x 36 */
    double* bp;          /* n_poses x 6 */
    double* Hll;         /* n_points x 9 */
    double* bl;          /* n_points x 3 */
    double* Hinv;        /* n_points x 9 (damped inverse of the last trial) */
    double* S;           /* n x n */
    double* bs;          /* n   */
    double* dp;          /* n_poses x 6 */
    double* dl;          /* n_points x 3 */
    int threads;
} work_t;

static void work_free(work_t* w)
{
    free(w->free_index);
    free(w->pt_ptr);
    free(w->perm);
    free(w->Rt);
    free(w->W);
    free(w->Hpp);
    free(w->bp);
    free(w->Hll);
    free(w->bl);
    free(w->Hinv);
    free(w->S);
    free(w->bs);
    free(w->dp);
    free(w->dl);
    memset(w, 0, sizeof(*w));
}

static int work_init(work_t* w, int32_t n_poses, const uint8_t* pose_fixed, int32_t n_points,
                     int64_t n_obs, const int32_t* obs_pose, const int32_t* obs_point, int threads)
{
    memset(w, 0, sizeof(*w));
    w->n_poses = n_poses;
    w->n_points = n_points;
    w->n_obs = n_obs;
    w->threads = threads < 1 ? 1 : threads;
    for (int64_t e = 0; e < n_obs; ++e) {
        if (obs_pose[e] < 0 || obs_pose[e] >= n_poses) return -2;
        if (obs_point[e] < 0 || obs_point[e] >= n_points) return -2;
    }
    w->free_index = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_poses > 0 ? n_poses : 1));
    int32_t nf = 0;
    for (int32_t c = 0; c < n_poses; ++c)
        w->free_index[c] = (pose_fixed && pose_fixed[c]) ? -1 : nf++;
    w->n_free = nf;
    w->n = 6 * nf;
    /* stable counting sort by point */
    w->pt_ptr = (int64_t*)calloc((size_t)n_points + 2, sizeof(int64_t));
    w->perm = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_obs > 0 ? n_obs : 1));
    for (int64_t e = 0; e < n_obs; ++e) w->pt_ptr[obs_point[e] + 1]++;
    for (int32_t p = 0; p < n_points; ++p) w->pt_ptr[p + 1] += w->pt_ptr[p];
    {
        int64_t* cur = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_points > 0 ? n_points : 1));
        for (int32_t p = 0; p < n_points; ++p) cur[p] = w->pt_ptr[p];
        for (int64_t e = 0; e < n_obs; ++e) w->perm[cur[obs_point[e]]++] = e;
        free(cur);
    }
    size_t nn = (size_t)w->n * (size_t)w->n;
    w->Rt = (double*)malloc(sizeof(double) * 12 * (size_t)(n_poses > 0 ? n_poses : 1));
    w->W = (double*)malloc(sizeof(double) * 18 * (size_t)(n_obs > 0 ? n_obs : 1));
    w->Hpp = (double*)malloc(sizeof(double) * 36 * (size_t)(n_poses > 0 ? n_poses : 1));
    w->bp = (double*)malloc(sizeof(double) * 6 * (size_t)(n_poses > 0 ? n_poses : 1));
    w->Hll = (double*)malloc(sizeof(double) * 9 * (size_t)(n_points > 0 ? n_points : 1));
    w->bl = (double*)malloc(sizeof(double) * 3 * (size_t)(n_points > 0 ? n_points : 1));
    w->Hinv = (double*)malloc(sizeof(double) * 9 * (size_t)(n_points > 0 ? n_points : 1));
    w->S = (double*)malloc(sizeof(double) * (nn > 0 ? nn : 1));
    w->bs = (double*)malloc(sizeof(double) * (size_t)(w->n > 0 ? w->n : 1));
    w->dp = (double*)calloc(6 * (size_t)(n_poses > 0 ? n_poses : 1), sizeof(double));
    w->dl = (double*)calloc(3 * (size_t)(n_points > 0 ? n_points : 1), sizeof(double));
    if (!w->Rt || !w->W || !w->Hpp || !w->bp || !w->Hll || !w->bl || !w->Hinv || !w->S ||
        !w->bs || !w->dp || !w->dl) {
        work_free(w);
        return -3;
    }
    return 0;
}

static void fill_Rt(work_t* w, const double* poses)
{
    for (int32_t c = 0; c < w->n_poses; ++c) {
        oracle_quat_to_R(poses + 7 * c, w->Rt + 12 * c);
        w->Rt[12 * c + 9] = poses[7 * c + 4];
        w->Rt[12 * c + 10] = poses[7 * c + 5];
        w->Rt[12 * c + 11] = poses[7 * c + 6];
    }
}

double oracle_cost(const oracle_cfg* c, int32_t n_poses, const double* poses, int32_t n_points,
                   const double* points, int64_t n_obs, const int32_t* obs_pose,
                   const int32_t* obs_point, const float* obs_uv)
{
    (void)n_points;
    double* Rt = (double*)malloc(sizeof(double) * 12 * (size_t)(n_poses > 0 ? n_poses : 1));
    for (int32_t k = 0; k < n_poses; ++k) {
        oracle_quat_to_R(poses + 7 * k, Rt + 12 * k);
        Rt[12 * k + 9] = poses[7 * k + 4];
        Rt[12 * k + 10] = poses[7 * k + 5];
        Rt[12 * k + 11] = poses[7 * k + 6];
    }
    double total = 0.0;
    int threads = c->threads < 1 ? 1 : c->threads;
    (void)threads;
#pragma omp parallel for reduction(+ : total) num_threads(threads) schedule(static)
    for (int64_t e = 0; e < n_obs; ++e) {
        const double* rt = Rt + 12 * obs_pose[e];
        double Pc[3];
        transform(rt, rt + 9, points + 3 * (size_t)obs_point[e], Pc);
        const double ru = c->fx * (Pc[0] / Pc[2]) + c->cx - (double)obs_uv[2 * e];
        const double rv = c->fy * (Pc[1] / Pc[2]) + c->cy - (double)obs_uv[2 * e + 1];
        double wgt;
        total += oracle_huber(ru * ru + rv * rv, c->huber_delta, &wgt);
    }
    free(Rt);
    return total;
}

/*
 * g2o buildSystem: per edge, e = proj - obs, rho = Huber(|e|^2),
 *   H += rho' J^T J,  b += -rho' J^T e      (robustInformation first-order form)
 */
static void build_system(const oracle_cfg* c, work_t* w, const double* poses, const double* points,
                         const int32_t* obs_pose, const float* obs_uv)
{
    fill_Rt(w, poses);
    const int T = w->threads;
    const size_t hsz = 36 * (size_t)w->n_poses, bsz = 6 * (size_t)w->n_poses;
    double* Hloc = (double*)calloc((hsz + bsz) * (size_t)T, sizeof(double));
#pragma omp parallel num_threads(T)
    {
#ifdef _OPENMP
        const int tid = omp_get_thread_num();
#else
        const int tid = 0;
#endif
        double* Hp = Hloc + (hsz + bsz) * (size_t)tid;
        double* Bp = Hp + hsz;
#pragma omp for schedule(static)
        for (int32_t p = 0; p < w->n_points; ++p) {
            double Hl[9] = { 0 }, Bl[3] = { 0 };
            for (int64_t s = w->pt_ptr[p]; s < w->pt_ptr[p + 1]; ++s) {
                const int64_t e = w->perm[s];
                const int32_t cam = obs_pose[e];
                const double* rt = w->Rt + 12 * cam;
                double Pc[3], Jp[12], Jl[6];
                transform(rt, rt + 9, points + 3 * (size_t)p, Pc);
                const double ru = c->fx * (Pc[0] / Pc[2]) + c->cx - (double)obs_uv[2 * e];
                const double rv = c->fy * (Pc[1] / Pc[2]) + c->cy - (double)obs_uv[2 * e + 1];
                double wgt;
                (void)oracle_huber(ru * ru + rv * rv, c->huber_delta, &wgt);
                oracle_jac_pose(c, Pc, Jp);
                oracle_jac_point(c, Pc, rt, Jl);
                for (int i = 0; i < 3; ++i) {
                    for (int j = 0; j < 3; ++j)
                        Hl[3 * i + j] += wgt * (Jl[i] * Jl[j] + Jl[3 + i] * Jl[3 + j]);
                    Bl[i] -= wgt * (Jl[i] * ru + Jl[3 + i] * rv);
                }
                double* We = w->W + 18 * s;
                if (w->free_index[cam] >= 0) {
                    double* H = Hp + 36 * cam;
                    double* B = Bp + 6 * cam;
                    for (int i = 0; i < 6; ++i) {
                        for (int j = 0; j < 6; ++j)
                            H[6 * i + j] += wgt * (Jp[i] * Jp[j] + Jp[6 + i] * Jp[6 + j]);
                        B[i] -= wgt * (Jp[i] * ru + Jp[6 + i] * rv);
                        for (int j = 0; j < 3; ++j)
                            We[3 * i + j] = wgt * (Jp[i] * Jl[j] + Jp[6 + i] * Jl[3 + j]);
                    }
                } else {
                    for (int i = 0; i < 18; ++i) We[i] = 0.0;
                }
            }
            memcpy(w->Hll + 9 * (size_t)p, Hl, sizeof(Hl));
            memcpy(w->bl + 3 * (size_t)p, Bl, sizeof(Bl));
        }
    }
    memset(w->Hpp, 0, sizeof(double) * hsz);
    memset(w->bp, 0, sizeof(double) * bsz);
    for (int t = 0; t < T; ++t) {
        const double* Hp = Hloc + (hsz + bsz) * (size_t)t;
        for (size_t i = 0; i < hsz; ++i) w->Hpp[i] += Hp[i];
        for (size_t i = 0; i < bsz; ++i) w->bp[i] += Hp[hsz + i];
    }
    free(Hloc);
}

/* g2o computeLambdaInit: max |H_jj| over every non-fixed vertex (poses and points) */
static double max_diagonal(const work_t* w)
{
    double m = 0.0;
    for (int32_t c = 0; c < w->n_poses; ++c) {
        if (w->free_index[c] < 0) continue;
        for (int i = 0; i < 6; ++i) m = fmax(m, fabs(w->Hpp[36 * c + 7 * i]));
    }
    for (int32_t p = 0; p < w->n_points; ++p)
        for (int i = 0; i < 3; ++i) m = fmax(m, fabs(w->Hll[9 * (size_t)p + 4 * i]));
    return m;
}

/* closed-form inverse of a symmetric 3x3 (what Eigen does for fixed 3x3) */
static void inv3(const double* A, double* Ai)
{
    const double c00 = A[4] * A[8] - A[5] * A[7];
    const double c01 = A[5] * A[6] - A[3] * A[8];
    const double c02 = A[3] * A[7] - A[4] * A[6];
    const double det = A[0] * c00 + A[1] * c01 + A[2] * c02;
    const double id = 1.0 / det;
    Ai[0] = c00 * id;
    Ai[1] = (A[2] * A[7] - A[1] * A[8]) * id;
    Ai[2] = (A[1] * A[5] - A[2] * A[4]) * id;
    Ai[3] = c01 * id;
    Ai[4] = (A[0] * A[8] - A[2] * A[6]) * id;
    Ai[5] = (A[2] * A[3] - A[0] * A[5]) * id;
    Ai[6] = c02 * id;
    Ai[7] = (A[1] * A[6] - A[0] * A[7]) * id;
    Ai[8] = (A[0] * A[4] - A[1] * A[3]) * id;
}

/*
 * g2o BlockSolver<6,3>::solve with marginalised points:
 *   S  = (Hpp + lambda I) - sum_l Hpl (Hll + lambda I)^-1 Hpl^T
 *   bs = bp - sum_l Hpl (Hll + lambda I)^-1 bl
 */
static void schur_reduce(work_t* w, const int32_t* obs_pose, double lambda)
{
    const int32_t n = w->n;
    const size_t nn = (size_t)n * (size_t)n;
    int T = w->threads;
    if ((double)nn * 8.0 * (double)T > 4.0e9) T = 1;
    double* Sloc = NULL;
    if (T > 1) Sloc = (double*)calloc((nn + (size_t)n) * (size_t)(T - 1), sizeof(double));
    memset(w->S, 0, sizeof(double) * nn);
    memset(w->bs, 0, sizeof(double) * (size_t)n);
    for (int32_t c = 0; c < w->n_poses; ++c) {
        const int32_t f = w->free_index[c];
        if (f < 0) continue;
        for (int i = 0; i < 6; ++i) {
            for (int j = 0; j < 6; ++j)
                w->S[(size_t)(6 * f + i) * n + 6 * f + j] = w->Hpp[36 * c + 6 * i + j];
            w->S[(size_t)(6 * f + i) * n + 6 * f + i] += lambda;
            w->bs[6 * f + i] = w->bp[6 * c + i];
        }
    }
#pragma omp parallel num_threads(T)
    {
#ifdef _OPENMP
        const int tid = omp_get_thread_num();
#else
        const int tid = 0;
#endif
        double* S = tid == 0 ? w->S : Sloc + (nn + (size_t)n) * (size_t)(tid - 1);
        double* bs = tid == 0 ? w->bs : S + nn;
#pragma omp for schedule(static)
        for (int32_t p = 0; p < w->n_points; ++p) {
            double Hd[9];
            memcpy(Hd, w->Hll + 9 * (size_t)p, sizeof(Hd));
            Hd[0] += lambda;
            Hd[4] += lambda;
            Hd[8] += lambda;
            double* Hi = w->Hinv + 9 * (size_t)p;
            inv3(Hd, Hi);
            const double* Bl = w->bl + 3 * (size_t)p;
            const int64_t s0 = w->pt_ptr[p], s1 = w->pt_ptr[p + 1];
            for (int64_t a = s0; a < s1; ++a) {
                const int32_t fa = w->free_index[obs_pose[w->perm[a]]];
                if (fa < 0) continue;
                const double* Wa = w->W + 18 * a;
                double Y[18];
                for (int i = 0; i < 6; ++i)
                    for (int j = 0; j < 3; ++j)
                        Y[3 * i + j] = Wa[3 * i] * Hi[j] + Wa[3 * i + 1] * Hi[3 + j] +
                                       Wa[3 * i + 2] * Hi[6 + j];
                for (int i = 0; i < 6; ++i)
                    bs[6 * fa + i] -= Y[3 * i] * Bl[0] + Y[3 * i + 1] * Bl[1] + Y[3 * i + 2] * Bl[2];
                for (int64_t b = s0; b < s1; ++b) {
                    const int32_t fb = w->free_index[obs_pose[w->perm[b]]];
                    if (fb < 0) continue;
                    const double* Wb = w->W + 18 * b;
                    double* Sb = S + (size_t)(6 * fa) * n + 6 * fb;
                    for (int i = 0; i < 6; ++i)
                        for (int j = 0; j < 6; ++j)
                            Sb[(size_t)i * n + j] -= Y[3 * i] * Wb[3 * j] +
                                                     Y[3 * i + 1] * Wb[3 * j + 1] +
                                                     Y[3 * i + 2] * Wb[3 * j + 2];
                }
            }
        }
    }
    for (int t = 1; t < T; ++t) {
        const double* S = Sloc + (nn + (size_t)n) * (size_t)(t - 1);
        for (size_t i = 0; i < nn; ++i) w->S[i] += S[i];
        for (int32_t i = 0; i < n; ++i) w->bs[i] += S[nn + i];
    }
    free(Sloc);
}

/* g2o LinearSolverDense on the pose system: dense Cholesky, lower, row-major, in place */
int32_t oracle_cholesky_solve(int32_t n, double* A, double* b, int32_t threads)
{
    if (threads < 1) threads = 1;
    for (int32_t j = 0; j < n; ++j) {
        double* Lj = A + (size_t)j * n;
        double d = Lj[j];
        for (int32_t k = 0; k < j; ++k) d -= Lj[k] * Lj[k];
        if (!(d > 0.0) || !isfinite(d)) return 1;
        d = sqrt(d);
        Lj[j] = d;
        const double id = 1.0 / d;
#pragma omp parallel for num_threads(threads) schedule(static) if (n - j > 256 && threads > 1)
        for (int32_t i = j + 1; i < n; ++i) {
            double* Li = A + (size_t)i * n;
            double s = Li[j];
            for (int32_t k = 0; k < j; ++k) s -= Li[k] * Lj[k];
            Li[j] = s * id;
        }
    }
    for (int32_t i = 0; i < n; ++i) {
        const double* Li = A + (size_t)i * n;
        double s = b[i];
        for (int32_t k = 0; k < i; ++k) s -= Li[k] * b[k];
        b[i] = s / Li[i];
    }
    for (int32_t i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int32_t k = i + 1; k < n; ++k) s -= A[(size_t)k * n + i] * b[k];
        b[i] = s / A[(size_t)i * n + i];
    }
    return 0;
}

/* x_l = (Hll + lambda I)^-1 (bl - Hpl^T x_p) */
static void back_substitute(work_t* w, const int32_t* obs_pose)
{
#pragma omp parallel for num_threads(w->threads) schedule(static)
    for (int32_t p = 0; p < w->n_points; ++p) {
        double r[3] = { w->bl[3 * (size_t)p], w->bl[3 * (size_t)p + 1], w->bl[3 * (size_t)p + 2] };
        for (int64_t a = w->pt_ptr[p]; a < w->pt_ptr[p + 1]; ++a) {
            const int32_t cam = obs_pose[w->perm[a]];
            if (w->free_index[cam] < 0) continue;
            const double* Wa = w->W + 18 * a;
            const double* d = w->dp + 6 * cam;
            for (int j = 0; j < 3; ++j)
                for (int i = 0; i < 6; ++i) r[j] -= Wa[3 * i + j] * d[i];
        }
        const double* Hi = w->Hinv + 9 * (size_t)p;
        for (int i = 0; i < 3; ++i)
            w->dl[3 * (size_t)p + i] = Hi[3 * i] * r[0] + Hi[3 * i + 1] * r[1] + Hi[3 * i + 2] * r[2];
    }
}

int32_t oracle_reduced_system(const oracle_cfg* c, int32_t n_poses, const double* poses,
                              const uint8_t* pose_fixed, int32_t n_points, const double* points,
                              int64_t n_obs, const int32_t* obs_pose, const int32_t* obs_point,
                              const float* obs_uv, double lambda, double* S, double* bs,
                              double* max_diag, int32_t* free_index)
{
    work_t w;
    int rc = work_init(&w, n_poses, pose_fixed, n_points, n_obs, obs_pose, obs_point, c->threads);
    if (rc) return rc;
    build_system(c, &w, poses, points, obs_pose, obs_uv);
    if (max_diag) *max_diag = max_diagonal(&w);
    schur_reduce(&w, obs_pose, lambda);
    if (S) memcpy(S, w.S, sizeof(double) * (size_t)w.n * (size_t)w.n);
    if (bs) memcpy(bs, w.bs, sizeof(double) * (size_t)w.n);
    if (free_index) memcpy(free_index, w.free_index, sizeof(int32_t) * (size_t)n_poses);
    int32_t n = w.n;
    work_free(&w);
    return n;
}

/* g2o OptimizationAlgorithmLevenberg::solve, restated (see DESIGN.md "LM policy") */
int32_t oracle_optimize(const oracle_cfg* c, int32_t n_poses, double* poses,
                        const uint8_t* pose_fixed, int32_t n_points, double* points,
                        int64_t n_obs, const int32_t* obs_pose, const int32_t* obs_point,
                        const float* obs_uv, int32_t iterations, oracle_iter_stats* stats,
                        double* final_cost)
{
    work_t w;
    int rc = work_init(&w, n_poses, pose_fixed, n_points, n_obs, obs_pose, obs_point, c->threads);
    if (rc) return rc;
    double* poses_new = (double*)malloc(sizeof(double) * 7 * (size_t)(n_poses > 0 ? n_poses : 1));
    double* points_new = (double*)malloc(sizeof(double) * 3 * (size_t)(n_points > 0 ? n_points : 1));
    double* Swork = (double*)malloc(sizeof(double) * ((size_t)w.n * (size_t)w.n + 1));
    double lambda = 0.0, ni = 2.0;
    double chi = oracle_cost(c, n_poses, poses, n_points, points, n_obs, obs_pose, obs_point, obs_uv);
    int32_t it = 0;
    for (; it < iterations; ++it) {
        build_system(c, &w, poses, points, obs_pose, obs_uv);
        if (it == 0) {
            lambda = c->lambda_init > 0.0 ? c->lambda_init : c->tau * max_diagonal(&w);
            ni = 2.0;
        }
        const double chi_before = chi;
        double rho = 0.0;
        int32_t q = 0, accepted = 0;
        do {
            schur_reduce(&w, obs_pose, lambda);
            /* keep bs: the scale term needs b, the solve overwrites a copy */
            memcpy(Swork, w.S, sizeof(double) * (size_t)w.n * (size_t)w.n);
            double* x = (double*)malloc(sizeof(double) * (size_t)(w.n > 0 ? w.n : 1));
            memcpy(x, w.bs, sizeof(double) * (size_t)w.n);
            const int ok2 = oracle_cholesky_solve(w.n, Swork, x, w.threads) == 0;
            for (int32_t k = 0; k < n_poses; ++k) {
                const int32_t f = w.free_index[k];
                for (int i = 0; i < 6; ++i) w.dp[6 * k + i] = f >= 0 ? x[6 * f + i] : 0.0;
            }
            free(x);
            back_substitute(&w, obs_pose);
            /* candidate state: poses by exp(d)*T, points by P + d */
            memcpy(poses_new, poses, sizeof(double) * 7 * (size_t)n_poses);
            for (int32_t k = 0; k < n_poses; ++k)
                if (w.free_index[k] >= 0) oracle_pose_oplus(w.dp + 6 * k, poses_new + 7 * k);
            for (size_t i = 0; i < 3 * (size_t)n_points; ++i) points_new[i] = points[i] + w.dl[i];
            double chi_new = oracle_cost(c, n_poses, poses_new, n_points, points_new, n_obs,
                                         obs_pose, obs_point, obs_uv);
            if (!ok2) chi_new = DBL_MAX;
            /* computeScale: sum_j x_j (lambda x_j + b_j), b = (bp, bl) before reduction */
            double scale = 0.0;
            for (int32_t k = 0; k < n_poses; ++k) {
                if (w.free_index[k] < 0) continue;
                for (int i = 0; i < 6; ++i)
                    scale += w.dp[6 * k + i] * (lambda * w.dp[6 * k + i] + w.bp[6 * k + i]);
            }
            for (size_t i = 0; i < 3 * (size_t)n_points; ++i)
                scale += w.dl[i] * (lambda * w.dl[i] + w.bl[i]);
            scale += 1e-3;
            rho = (chi - chi_new) / scale;
            if (rho > 0.0 && isfinite(chi_new)) {
                double alpha = 1.0 - pow(2.0 * rho - 1.0, 3.0);
                alpha = fmin(alpha, 2.0 / 3.0);
                const double sf = fmax(1.0 / 3.0, alpha);
                lambda *= sf;
                ni = 2.0;
                chi = chi_new;
                memcpy(poses, poses_new, sizeof(double) * 7 * (size_t)n_poses);
                memcpy(points, points_new, sizeof(double) * 3 * (size_t)n_points);
                accepted = 1;
            } else {
                lambda *= ni;
                ni *= 2.0;
                if (!isfinite(lambda)) {
                    ++q;
                    break;
                }
            }
            ++q;
        } while (rho < 0.0 && q < c->max_trials);
        if (stats) {
            stats[it].cost_before = chi_before;
            stats[it].cost_after = chi;
            stats[it].lambda = lambda;
            stats[it].trials = q;
            stats[it].accepted = accepted;
        }
        if (q == c->max_trials || rho == 0.0 || !isfinite(lambda)) {
            ++it;
            break;
        }
    }
    if (final_cost) *final_cost = chi;
    free(poses_new);
    free(points_new);
    free(Swork);
    work_free(&w);
    return it;
}
