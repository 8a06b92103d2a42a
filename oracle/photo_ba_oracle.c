/*
 * photo_ba_oracle.c — CPU restatement of bundle adjustment over the reference adjuster's OWN edge,
 * EdgeDirectProjection (reference include/bundle_adjuster.h:58-106).  TEST INFRASTRUCTURE ONLY: only
 * tests/, __graft_entry__.smoke() and bench.py's CPU legs may load it.
 *
 * PARITY UNPINNED (as for ba_oracle.c): the reference never adds such an edge to a running optimiser
 * (bundle_adjuster.h:154-165 does not compile) and has no tests.  What is restated:
 *   - the 16-D residual of computeError (:70-101) statement by statement: target pixel K (R p + t) / z
 *     with the pose VERTEX (:73-78), source pixel srcFrame->Project(p) with the keyframe's own constant
 *     pose (:79, keyframe.h:82-89), the 4 x 4 patch i, j = -2..1 (:83-84), the float casts of the sample
 *     coordinates (:86-89), GetPixelValue (keyframe.h:50-57), information I_16 (:160);
 *   - an edge one of whose samples leaves an image disables itself (setLevel(1), :91-95): it
 *     contributes nothing (no residual, no Jacobian, no cost) AT THAT STATE;
 *   - the reference defines no linearizeOplus, so g2o would difference computeError numerically; here
 *     the Jacobian is analytic: image gradient (keyframe.h:59-64, central differences of the sampler)
 *     times dPixeldXi (viso.cpp:640-658) for the pose, times the 2 x 3 projection Jacobians of both
 *     frames for the point.  tests/test_photo_ba_oracle.py checks it against central differences of a
 *     smooth statement of the residual.
 *   - g2o's Levenberg-Marquardt policy and Huber kernel exactly as in ba_oracle.c::oracle_optimize.
 *
 * Deliberately written differently from the device code: explicit 16-row Jacobians per edge and ONE
 * dense normal-equation matrix over [free poses | points] with a plain Cholesky — no Schur
 * complement, no 2 x 2 gradient sums.
 */
#include "photo_ba_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

static int inside(const direct_image* im, double u, double v)
{ /* keyframe.h:76-79 */
    return u >= 0 && u < im->cols && v >= 0 && v < im->rows;
}

/* residual (16), validity, and the two 16 x 2 gradient stacks of one edge at the given state */
static int edge_terms(const oracle_cfg* c, const direct_image* src, const direct_image* tgt, const double src_pose[7],
                      const double dst_pose[7], const double p[3], double r[16], double gs[32], double gt[32],
                      double Pc_s[3], double Pc_t[3], double R_s[9], double R_t[9])
{
    oracle_quat_to_R(dst_pose, R_t);
    oracle_quat_to_R(src_pose, R_s);
    for (int i = 0; i < 3; ++i) {
        Pc_t[i] = R_t[3 * i] * p[0] + R_t[3 * i + 1] * p[1] + R_t[3 * i + 2] * p[2] + dst_pose[4 + i];
        Pc_s[i] = R_s[3 * i] * p[0] + R_s[3 * i + 1] * p[1] + R_s[3 * i + 2] * p[2] + src_pose[4 + i];
    }
    /* :77-78  target_uv1 = K * (R p + t); target_uv1 /= z */
    const double tu = (c->fx * Pc_t[0] + 0.0 * Pc_t[1] + c->cx * Pc_t[2]) / Pc_t[2];
    const double tv = (0.0 * Pc_t[0] + c->fy * Pc_t[1] + c->cy * Pc_t[2]) / Pc_t[2];
    /* :79, keyframe.h:82-89 at level 0 */
    const double su = 1.0 * ((Pc_s[0] / Pc_s[2]) * c->fx + c->cx), sv = 1.0 * ((Pc_s[1] / Pc_s[2]) * c->fy + c->cy);
    int idx = 0;
    for (int j = -2; j <= 1; ++j)
        for (int i = -2; i <= 1; ++i, ++idx) {
            const float u1 = (float)(tu + (double)i), v1 = (float)(tv + (double)j); /* :86-87 */
            const float u2 = (float)(su + (double)i), v2 = (float)(sv + (double)j); /* :88-89 */
            if (!inside(tgt, u1, v1) || !inside(src, u2, v2)) return 0;            /* :91-95 */
            r[idx] = direct_pixel_value(src, u2, v2) - direct_pixel_value(tgt, u1, v1); /* :97 */
            if (gs) direct_gradient(src, u2, v2, gs + 2 * idx);
            if (gt) direct_gradient(tgt, u1, v1, gt + 2 * idx);
        }
    return 1;
}

double photo_ba_cost(const oracle_cfg* c, const direct_image* images, int32_t n_poses, const double* poses,
                     const double* src_poses, const double* points, int64_t n_edges, const int32_t* e_point,
                     const int32_t* e_src, const int32_t* e_dst)
{
    (void)n_poses;
    double chi = 0.0;
    for (int64_t e = 0; e < n_edges; ++e) {
        double r[16], Ps[3], Pt[3], Rs[9], Rt[9];
        if (!edge_terms(c, images + e_src[e], images + e_dst[e], src_poses + 7 * (size_t)e_src[e], poses + 7 * (size_t)e_dst[e],
                        points + 3 * (size_t)e_point[e], r, NULL, NULL, Ps, Pt, Rs, Rt))
            continue;
        double s = 0.0, w;
        for (int k = 0; k < 16; ++k) s += r[k] * r[k];
        chi += oracle_huber(s, c->huber_delta, &w);
    }
    return chi;
}

int32_t photo_ba_edge_jacobian(const oracle_cfg* c, const direct_image* src, const direct_image* tgt, const double src_pose[7],
                               const double dst_pose[7], const double p[3], double r[16], double Jp[96], double Jl[48])
{
    double gs[32], gt[32], Ps[3], Pt[3], Rs[9], Rt[9], B[12], As[6], At[6];
    if (!edge_terms(c, src, tgt, src_pose, dst_pose, p, r, gs, gt, Ps, Pt, Rs, Rt)) return 0;
    oracle_jac_pose(c, Pt, B);        /* viso.cpp:640-658 */
    oracle_jac_point(c, Pt, Rt, At);  /* first three columns of dPixeldXi times R */
    oracle_jac_point(c, Ps, Rs, As);
    for (int k = 0; k < 16; ++k) {
        for (int a = 0; a < 6; ++a) Jp[6 * k + a] = -(gt[2 * k] * B[a] + gt[2 * k + 1] * B[6 + a]);
        for (int a = 0; a < 3; ++a)
            Jl[3 * k + a] = (gs[2 * k] * As[a] + gs[2 * k + 1] * As[3 + a]) - (gt[2 * k] * At[a] + gt[2 * k + 1] * At[3 + a]);
    }
    return 1;
}

/* dense H (n x n, row-major, lower and upper filled) and b = -J^T W r over [free poses | points]; returns max diag(H) */
static double build_dense(const oracle_cfg* c, const direct_image* images, const double* poses, const double* src_poses,
                          const int32_t* free_index, int32_t nf, int32_t n_points, const double* points, int64_t n_edges,
                          const int32_t* e_point, const int32_t* e_src, const int32_t* e_dst, double* H, double* b)
{
    const int32_t n = 6 * nf + 3 * n_points;
    memset(H, 0, sizeof(double) * (size_t)n * (size_t)n);
    memset(b, 0, sizeof(double) * (size_t)n);
    for (int64_t e = 0; e < n_edges; ++e) {
        double r[16], Jp[96], Jl[48];
        const int32_t d = e_dst[e], l = e_point[e];
        if (!photo_ba_edge_jacobian(c, images + e_src[e], images + d, src_poses + 7 * (size_t)e_src[e], poses + 7 * (size_t)d,
                                    points + 3 * (size_t)l, r, Jp, Jl))
            continue;
        double s = 0.0, w;
        for (int k = 0; k < 16; ++k) s += r[k] * r[k];
        oracle_huber(s, c->huber_delta, &w);
        const int32_t f = free_index[d];
        const int32_t op = 6 * f, ol = 6 * nf + 3 * l;
        for (int k = 0; k < 16; ++k) {
            const double* jp = Jp + 6 * k;
            const double* jl = Jl + 3 * k;
            for (int a = 0; a < 3; ++a) {
                b[ol + a] -= w * jl[a] * r[k];
                for (int q = 0; q < 3; ++q) H[(size_t)(ol + a) * n + ol + q] += w * jl[a] * jl[q];
            }
            if (f < 0) continue;
            for (int a = 0; a < 6; ++a) {
                b[op + a] -= w * jp[a] * r[k];
                for (int q = 0; q < 6; ++q) H[(size_t)(op + a) * n + op + q] += w * jp[a] * jp[q];
                for (int q = 0; q < 3; ++q) {
                    H[(size_t)(op + a) * n + ol + q] += w * jp[a] * jl[q];
                    H[(size_t)(ol + q) * n + op + a] += w * jp[a] * jl[q];
                }
            }
        }
    }
    double m = 0.0;
    for (int32_t i = 0; i < n; ++i) m = fmax(m, fabs(H[(size_t)i * n + i]));
    return m;
}

int32_t photo_ba_optimize(const oracle_cfg* c, const direct_image* images, int32_t n_poses, double* poses,
                          const uint8_t* pose_fixed, int32_t n_points, double* points, int64_t n_edges,
                          const int32_t* e_point, const int32_t* e_src, const int32_t* e_dst, int32_t iterations,
                          oracle_iter_stats* stats, double* final_cost)
{
    if (n_poses <= 0 || n_points < 0 || n_edges < 0) return -1;
    int32_t* free_index = (int32_t*)malloc(sizeof(int32_t) * (size_t)n_poses);
    int32_t nf = 0;
    for (int32_t k = 0; k < n_poses; ++k) free_index[k] = (pose_fixed && pose_fixed[k]) ? -1 : nf++;
    const int32_t n = 6 * nf + 3 * n_points;
    /* srcFrame->Project uses the Keyframe's own R_, T_ (keyframe.h:82-89), which the optimiser never touches:
       the source poses stay at the values the keyframes had when the edges were created */
    double* src_poses = (double*)malloc(sizeof(double) * 7 * (size_t)n_poses);
    memcpy(src_poses, poses, sizeof(double) * 7 * (size_t)n_poses);
    double* H = (double*)malloc(sizeof(double) * ((size_t)n * (size_t)n + 1));
    double* Hd = (double*)malloc(sizeof(double) * ((size_t)n * (size_t)n + 1));
    double* b = (double*)malloc(sizeof(double) * (size_t)(n + 1));
    double* x = (double*)malloc(sizeof(double) * (size_t)(n + 1));
    double* poses_new = (double*)malloc(sizeof(double) * 7 * (size_t)n_poses);
    double* points_new = (double*)malloc(sizeof(double) * 3 * (size_t)(n_points > 0 ? n_points : 1));
    double lambda = 0.0, ni = 2.0;
    double chi = photo_ba_cost(c, images, n_poses, poses, src_poses, points, n_edges, e_point, e_src, e_dst);
    int32_t it = 0;
    for (; it < iterations; ++it) {
        const double maxd = build_dense(c, images, poses, src_poses, free_index, nf, n_points, points, n_edges, e_point, e_src, e_dst, H, b);
        if (it == 0) {
            lambda = c->lambda_init > 0.0 ? c->lambda_init : c->tau * maxd;
            ni = 2.0;
        }
        const double chi_before = chi;
        double rho = 0.0;
        int32_t q = 0, accepted = 0;
        do {
            memcpy(Hd, H, sizeof(double) * (size_t)n * (size_t)n);
            for (int32_t i = 0; i < n; ++i) Hd[(size_t)i * n + i] += lambda;
            memcpy(x, b, sizeof(double) * (size_t)n);
            const int ok2 = oracle_cholesky_solve(n, Hd, x, c->threads) == 0;
            memcpy(poses_new, poses, sizeof(double) * 7 * (size_t)n_poses);
            for (int32_t k = 0; k < n_poses; ++k)
                if (free_index[k] >= 0) oracle_pose_oplus(x + 6 * free_index[k], poses_new + 7 * k);
            for (size_t i = 0; i < 3 * (size_t)n_points; ++i) points_new[i] = points[i] + x[6 * (size_t)nf + i];
            double chi_new = photo_ba_cost(c, images, n_poses, poses_new, src_poses, points_new, n_edges, e_point, e_src, e_dst);
            if (!ok2) chi_new = DBL_MAX;
            double scale = 0.0;
            for (int32_t i = 0; i < n; ++i) scale += x[i] * (lambda * x[i] + b[i]);
            scale += 1e-3;
            rho = (chi - chi_new) / scale;
            if (rho > 0.0 && isfinite(chi_new)) {
                double alpha = 1.0 - pow(2.0 * rho - 1.0, 3.0);
                alpha = fmin(alpha, 2.0 / 3.0);
                lambda *= fmax(1.0 / 3.0, alpha);
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
    free(free_index), free(src_poses), free(H), free(Hd), free(b), free(x), free(poses_new), free(points_new);
    return it;
}
