/*
 * direct_oracle.c — CPU restatement of viso's direct pose estimation.  TEST INFRASTRUCTURE ONLY
 * (see direct_oracle.h).  Plain C11, fp64, sequential sums in the reference's loop order.
 * Citations are relative to /root/reference.
 */
#include "direct_oracle.h"
#include "ba_oracle.h" /* oracle_quat_to_R, oracle_pose_oplus (Sophus SE3d::exp(xi) * T) */

#include <math.h>
#include <string.h>

void direct_default_cfg(direct_cfg* c)
{
    memset(c, 0, sizeof(*c));
    c->fx = 517.3, c->fy = 516.5, c->cx = 325.1, c->cy = 249.7; /* src/main.cpp:14-17 */
    c->half_patch_size = 4;                                     /* include/viso.h:25 */
    c->max_iterations = 100;                                    /* src/viso.cpp:671 */
    c->delta_thresh = 0.005;                                    /* src/viso.cpp:668 */
    c->reference_termination = 1;
}

/* One byte of the cv::Mat buffer.  The reference indexes data[...] without a bounds check; where
 * that leaves the buffer (undefined behaviour there) the address is clamped into it. */
static double tap(const direct_image* im, long idx)
{
    const long last = (long)im->rows * im->step - 1;
    if (idx < 0) idx = 0;
    if (idx > last) idx = last;
    return (double)im->data[idx];
}

/* include/keyframe.h:50-57  GetPixelValue: int() truncation for the address, floor() for the weights */
double direct_pixel_value(const direct_image* im, double x, double y)
{
    const long base = (long)(int)y * im->step + (long)(int)x;
    const double xx = x - floor(x);
    const double yy = y - floor(y);
    return (1 - xx) * (1 - yy) * tap(im, base) + xx * (1 - yy) * tap(im, base + 1)
        + (1 - xx) * yy * tap(im, base + im->step) + xx * yy * tap(im, base + im->step + 1);
}

/* include/keyframe.h:59-64  GetGradient */
void direct_gradient(const direct_image* im, double u, double v, double g[2])
{
    g[0] = 0.5 * (direct_pixel_value(im, u + 1, v) - direct_pixel_value(im, u - 1, v));
    g[1] = 0.5 * (direct_pixel_value(im, u, v + 1) - direct_pixel_value(im, u, v - 1));
}

/* include/keyframe.h:76-79  IsInside(u, v, level) */
static int is_inside(const direct_image* im, double u, double v)
{
    return u >= 0 && u < im->cols && v >= 0 && v < im->rows;
}

/* include/keyframe.h:82-89  Project(point, level): uv1 = R P + T; uv1 /= z; scale * (x fx + cx) */
void direct_project(const direct_cfg* c, const double pose[7], const double Pw[3], double scale, double uv[2])
{
    double R[9];
    oracle_quat_to_R(pose, R);
    double x = R[0] * Pw[0] + R[1] * Pw[1] + R[2] * Pw[2] + pose[4];
    double y = R[3] * Pw[0] + R[4] * Pw[1] + R[5] * Pw[2] + pose[5];
    const double z = R[6] * Pw[0] + R[7] * Pw[1] + R[8] * Pw[2] + pose[6];
    x /= z;
    y /= z;
    uv[0] = scale * (x * c->fx + c->cx);
    uv[1] = scale * (y * c->fy + c->cy);
}

/* src/viso.cpp:640-658  dPixeldXi, row-major 2x6 */
void direct_dpixel_dxi(const direct_cfg* c, const double R[9], const double T[3], const double P[3],
                       double scale, double J[12])
{
    const double x = R[0] * P[0] + R[1] * P[1] + R[2] * P[2] + T[0];
    const double y = R[3] * P[0] + R[4] * P[1] + R[5] * P[2] + T[1];
    const double z = R[6] * P[0] + R[7] * P[1] + R[8] * P[2] + T[2];
    const double fx = c->fx * scale, fy = c->fy * scale;
    const double zz = z * z, xy = x * y;
    J[0] = fx / z, J[1] = 0, J[2] = -fx * x / zz, J[3] = -fx * xy / zz, J[4] = fx + fx * x * x / zz, J[5] = -fx * y / z;
    J[6] = 0, J[7] = fy / z, J[8] = -fy * y / zz, J[9] = -fy - fy * y * y / zz, J[10] = fy * xy / zz, J[11] = fy * x / z;
}

/* src/viso.cpp:688-733: the loop over the map points.  `cost` is ADDED to (the caller decides
 * whether it starts from zero); H, b and n_good start from zero (viso.cpp:678-683). */
static void accumulate(const direct_cfg* c, int32_t level, const direct_image* ref, const direct_image* cur,
                       const double ref_pose[7], const double T21[7], const double* points, int32_t n_points,
                       double H[36], double b[6], double* cost, int32_t* n_good)
{
    static const double scales[] = { 1, 0.5, 0.25, 0.125 }; /* viso.cpp:666 */
    const double scale = scales[level];
    const double hp = (double)c->half_patch_size;
    double R[9];
    oracle_quat_to_R(T21, R); /* current_frame->SetR(T21.rotationMatrix()), viso.cpp:685-686 */
    memset(H, 0, 36 * sizeof(double));
    memset(b, 0, 6 * sizeof(double));
    *n_good = 0;
    for (int32_t i = 0; i < n_points; ++i) {
        const double* P1 = points + 3 * (size_t)i;
        double uv_ref[2], uv_cur[2];
        direct_project(c, ref_pose, P1, scale, uv_ref); /* viso.cpp:693-695 */
        direct_project(c, T21, P1, scale, uv_cur);      /* viso.cpp:697-699 */
        const int good = is_inside(ref, uv_ref[0] - hp, uv_ref[1] - hp) && is_inside(ref, uv_ref[0] + hp, uv_ref[1] + hp)
            && is_inside(cur, uv_cur[0] - hp, uv_cur[1] - hp) && is_inside(cur, uv_cur[0] + hp, uv_cur[1] + hp);
        if (!good) continue; /* viso.cpp:704-715 (NaN projections fail every comparison) */
        ++*n_good;
        double Jx[12];
        direct_dpixel_dxi(c, R, T21 + 4, P1, scale, Jx); /* viso.cpp:719-720 */
        for (int x = -c->half_patch_size; x < c->half_patch_size; ++x) {
            for (int y = -c->half_patch_size; y < c->half_patch_size; ++y) {
                const double error = direct_pixel_value(ref, uv_ref[0] + x, uv_ref[1] + y)
                    - direct_pixel_value(cur, uv_cur[0] + x, uv_cur[1] + y);
                double g[2];
                direct_gradient(cur, uv_cur[0] + x, uv_cur[1] + y, g);
                double J[6];
                for (int k = 0; k < 6; ++k) J[k] = -(g[0] * Jx[k] + g[1] * Jx[6 + k]); /* viso.cpp:726 */
                for (int r = 0; r < 6; ++r)
                    for (int q = 0; q < 6; ++q) H[6 * r + q] += J[r] * J[q]; /* viso.cpp:727 */
                for (int k = 0; k < 6; ++k) b[k] += -error * J[k];            /* viso.cpp:728 */
                *cost += error * error;                                       /* viso.cpp:729 */
            }
        }
    }
}

void direct_linearize(const direct_cfg* c, int32_t level, const direct_image* ref, const direct_image* cur,
                      const double ref_pose[7], const double T21[7], const double* points, int32_t n_points,
                      double H[36], double b[6], double* cost, int32_t* n_good)
{
    *cost = 0.0;
    accumulate(c, level, ref, cur, ref_pose, T21, points, n_points, H, b, cost, n_good);
}

/* update = H.inverse() * b (viso.cpp:735).  Eigen's inverse() of a dynamic-free 6x6 goes through PartialPivLU: P H = L U with
 * row pivoting, then the inverse column by column from the unit vectors.  Restated exactly that way (Doolittle LU in
 * place, forward / backward substitution per column of the identity, then the product with b) — deliberately NOT the
 * Gauss-Jordan tableau the device kernel uses (viso_b200/csrc/direct_pose.cu), so that the two statements of this solve
 * are independent texts. */
static void inverse_times(const double H[36], const double b[6], double x[6])
{
    double LU[6][6], inv[6][6];
    int perm[6];
    for (int r = 0; r < 6; ++r) {
        perm[r] = r;
        for (int q = 0; q < 6; ++q) LU[r][q] = H[6 * r + q];
    }
    for (int k = 0; k < 6; ++k) { /* P H = L U: unit lower L below the diagonal, U on and above it */
        int p = k;
        for (int r = k + 1; r < 6; ++r)
            if (fabs(LU[r][k]) > fabs(LU[p][k])) p = r;
        if (p != k) {
            for (int q = 0; q < 6; ++q) {
                const double t = LU[k][q];
                LU[k][q] = LU[p][q], LU[p][q] = t;
            }
            const int t = perm[k];
            perm[k] = perm[p], perm[p] = t;
        }
        for (int r = k + 1; r < 6; ++r) { /* singular H: inf / NaN propagate like in the reference */
            LU[r][k] /= LU[k][k];
            for (int q = k + 1; q < 6; ++q) LU[r][q] -= LU[r][k] * LU[k][q];
        }
    }
    for (int col = 0; col < 6; ++col) { /* column `col` of the inverse: L U z = P e_col */
        double y[6];
        for (int r = 0; r < 6; ++r) {
            double v = perm[r] == col ? 1.0 : 0.0;
            for (int q = 0; q < r; ++q) v -= LU[r][q] * y[q];
            y[r] = v;
        }
        for (int r = 5; r >= 0; --r) {
            double v = y[r];
            for (int q = r + 1; q < 6; ++q) v -= LU[r][q] * inv[q][col];
            inv[r][col] = v / LU[r][r];
        }
    }
    for (int r = 0; r < 6; ++r) {
        double v = 0.0;
        for (int q = 0; q < 6; ++q) v += inv[r][q] * b[q];
        x[r] = v;
    }
}

/* src/viso.cpp:661-758 */
void direct_single_layer(const direct_cfg* c, int32_t level, const direct_image* ref, const direct_image* cur,
                         const double ref_pose[7], double T21[7], const double* points, int32_t n_points,
                         direct_level_stats* stats)
{
    double cost = 0, lastCost = 0; /* viso.cpp:673 */
    int32_t nGood = 0;
    double best[7];
    memcpy(best, T21, sizeof(best)); /* viso.cpp:676 */
    int32_t applied = 0, reason = 0;
    for (int iter = 0; iter < c->max_iterations; ++iter) {
        double H[36], b[6], update[6];
        if (!c->reference_termination) cost = 0; /* the reference never resets it */
        accumulate(c, level, ref, cur, ref_pose, T21, points, n_points, H, b, &cost, &nGood);
        inverse_times(H, b, update);     /* viso.cpp:735 */
        oracle_pose_oplus(update, T21);  /* viso.cpp:737 */
        ++applied;
        cost /= nGood;                   /* viso.cpp:739 */
        if (isnan(update[0])) {          /* viso.cpp:741-744 */
            memcpy(T21, best, sizeof(best));
            reason = 1;
            break;
        }
        if (iter > 0 && cost > lastCost) { /* viso.cpp:746-749 */
            memcpy(T21, best, sizeof(best));
            reason = 2;
            break;
        }
        if (c->reference_termination || iter > 0) {
            if ((1 - cost / (double)lastCost) < c->delta_thresh) { /* viso.cpp:751-753 */
                reason = 3;
                break;
            }
        }
        memcpy(best, T21, sizeof(best)); /* viso.cpp:755-756 */
        lastCost = cost;
    }
    if (stats) {
        stats->iterations = applied;
        stats->n_good = nGood;
        stats->stop_reason = reason;
        stats->reserved = 0;
        stats->cost = cost;
        stats->last_cost = lastCost;
    }
}

/* src/viso.cpp:760-766 */
void direct_multi_layer(const direct_cfg* c, const direct_image* ref, const direct_image* cur,
                        const double ref_pose[7], double T21[7], const double* points, int32_t n_points,
                        direct_level_stats* stats)
{
    for (int level = 3; level >= 0; --level)
        direct_single_layer(c, level, ref + level, cur + level, ref_pose, T21, points, n_points,
            stats ? stats + level : 0);
}
