/*
 * lk_oracle.c — CPU restatement of viso's Lucas-Kanade patch alignment.  TEST INFRASTRUCTURE ONLY
 * (see lk_oracle.h).  Plain C11; sums in the reference's loop order; float where the reference
 * holds cv::Point2f.  Citations are relative to /root/reference.
 */
#include "lk_oracle.h"
#include "ba_oracle.h" /* oracle_quat_to_R */

#include <math.h>
#include <stdlib.h>
#include <string.h>

static const double kScales[4] = { 1.0, 0.5, 0.25, 0.125 }; /* viso.cpp:364, :850 */

void lk_default_cfg(lk_cfg* c)
{
    memset(c, 0, sizeof(*c));
    c->half_patch_size = 4;
    c->align_iterations = 100;
    c->flow_iterations = 10;
    c->photometric_error_thresh = (4 * 2) * (4 * 2) * 15 * 15; /* include/viso.h:26 */
}

static int is_inside(const direct_image* im, double u, double v) /* keyframe.h:76-79 */
{
    return u >= 0 && u < im->cols && v >= 0 && v < im->rows;
}

/* update = H.inverse() * b for a 2x2 H: Eigen's closed form (1/det times the adjugate) */
static void solve2(const double H[4], const double b[2], double u[2])
{
    const double invdet = 1.0 / (H[0] * H[3] - H[1] * H[2]);
    const double i00 = H[3] * invdet, i10 = -H[2] * invdet, i01 = -H[1] * invdet, i11 = H[0] * invdet;
    u[0] = i00 * b[0] + i01 * b[1];
    u[1] = i10 * b[0] + i11 * b[1];
}

/* viso.cpp:845-925 */
void lk_align_single(const lk_cfg* c, const direct_image* ref_images, const int32_t* ref_slot, const direct_image* cur_pyr,
                     int32_t n, const double* uv_ref, double* uv_cur, uint8_t* success, int32_t level)
{
    const double s = kScales[level], hp = (double)c->half_patch_size;
    const direct_image* cur = cur_pyr + level;
    for (int32_t i = 0; i < n; ++i) {
        const direct_image* ref = ref_images + 4 * (size_t)ref_slot[i] + level;
        const double urx = uv_ref[2 * i], ury = uv_ref[2 * i + 1];
        const double ucx = uv_cur[2 * i], ucy = uv_cur[2 * i + 1];
        double dx = 0, dy = 0, cost = 0, lastCost = 0; /* viso.cpp:858-860 */
        int succ = 1;
        for (int iter = 0; iter < c->align_iterations; ++iter) {
            double H[4] = { 0, 0, 0, 0 }, b[2] = { 0, 0 };
            cost = 0;
            if (!is_inside(ref, urx * s + dx - hp, ury * s + dy - hp) || !is_inside(ref, urx * s + dx + hp, ury * s + dy + hp)) {
                succ = 0; /* viso.cpp:869-873 */
                break;
            }
            for (int x = -c->half_patch_size; x < c->half_patch_size; ++x)
                for (int y = -c->half_patch_size; y < c->half_patch_size; ++y) {
                    double g[2];
                    direct_gradient(ref, urx * s + x, ury * s + y, g); /* inverse = true, viso.cpp:882-883 */
                    const double J0 = -g[0], J1 = -g[1];
                    const double error = direct_pixel_value(ref, urx * s + x, ury * s + y)
                        - direct_pixel_value(cur, ucx * s + x + dx, ucy * s + y + dy); /* viso.cpp:885 */
                    H[0] += J0 * J0, H[1] += J0 * J1, H[2] += J1 * J0, H[3] += J1 * J1; /* viso.cpp:888 */
                    b[0] += -J0 * error, b[1] += -J1 * error;
                    cost += error * error;
                }
            double update[2];
            solve2(H, b, update); /* viso.cpp:894 */
            if (isnan(update[0])) { /* viso.cpp:896-899 */
                succ = 0;
                break;
            }
            if (iter > 0 && cost > lastCost) break; /* viso.cpp:900-902 */
            dx += update[0], dy += update[1];
            lastCost = cost;
            succ = !(lastCost > c->photometric_error_thresh); /* viso.cpp:909-913 */
        }
        success[i] = (uint8_t)succ;
        uv_cur[2 * i] = ucx + dx / s, uv_cur[2 * i + 1] = ucy + dy / s; /* viso.cpp:916 */
    }
}

/* viso.cpp:828-830: success of the last (finest) level is what the caller keeps */
void lk_align(const lk_cfg* c, const direct_image* ref_images, const int32_t* ref_slot, const direct_image* cur,
              int32_t n, const double* uv_ref, double* uv_cur, uint8_t* success)
{
    for (int level = 3; level >= 0; --level) lk_align_single(c, ref_images, ref_slot, cur, n, uv_ref, uv_cur, success, level);
}

/* viso.cpp:259-351.  kp.pt is cv::Point2f: `kp.pt.x + x` is a float sum, `+ dx` promotes to double. */
void lk_flow_single(const lk_cfg* c, const direct_image* img1, const direct_image* img2, int32_t n,
                    const float* kp1, float* kp2, int32_t have_initial, int32_t inverse, uint8_t* success)
{
    const double hp = (double)c->half_patch_size;
    for (int32_t i = 0; i < n; ++i) {
        const float kx = kp1[2 * i], ky = kp1[2 * i + 1];
        double dx = 0, dy = 0;
        if (have_initial) { /* viso.cpp:272-275: float differences */
            dx = (double)(float)(kp2[2 * i] - kx);
            dy = (double)(float)(kp2[2 * i + 1] - ky);
        }
        double cost = 0, lastCost = 0;
        int succ = 1;
        for (int iter = 0; iter < c->flow_iterations; ++iter) {
            double H[4] = { 0, 0, 0, 0 }, b[2] = { 0, 0 };
            cost = 0;
            if (kx + dx <= hp || kx + dx >= img1->cols - hp || ky + dy <= hp || ky + dy >= img1->rows - hp) {
                succ = 0; /* viso.cpp:286-289 */
                break;
            }
            for (int x = -c->half_patch_size; x < c->half_patch_size; ++x)
                for (int y = -c->half_patch_size; y < c->half_patch_size; ++y) {
                    const float fx = kx + (float)x, fy = ky + (float)y; /* float + int -> float */
                    double g[2];
                    if (!inverse) direct_gradient(img2, fx + dx, fy + dy, g); /* viso.cpp:299 */
                    else direct_gradient(img1, fx, fy, g);                    /* viso.cpp:302 */
                    const double J0 = -g[0], J1 = -g[1];
                    const double error = direct_pixel_value(img1, fx, fy) - direct_pixel_value(img2, fx + dx, fy + dy); /* :305 */
                    H[0] += J0 * J0, H[1] += J0 * J1, H[2] += J1 * J0, H[3] += J1 * J1;
                    b[0] += -J0 * error, b[1] += -J1 * error;
                    cost += error * error;
                }
            double update[2];
            solve2(H, b, update); /* viso.cpp:313 */
            if (isnan(update[0])) { /* viso.cpp:315-321 */
                succ = 0;
                break;
            }
            if (iter > 0 && cost > lastCost) break; /* viso.cpp:323-325 */
            dx += update[0], dy += update[1];
            lastCost = cost;
            succ = !(lastCost > c->photometric_error_thresh); /* viso.cpp:332-336 */
        }
        success[i] = (uint8_t)succ;
        kp2[2 * i] = kx + (float)dx, kp2[2 * i + 1] = ky + (float)dy; /* viso.cpp:342-349: Point2f sum */
    }
}

/* viso.cpp:353-391.  cv::Point2f *= double and /= double round the double product to float. */
void lk_flow_multi(const lk_cfg* c, const direct_image* pyr1, const direct_image* pyr2, int32_t n,
                   const float* kp1, float* kp2, int32_t have_initial, int32_t inverse, uint8_t* success)
{
    const double pyramid_scale = 0.5;
    float* kp1_ = (float*)malloc(sizeof(float) * 2 * (size_t)(n > 0 ? n : 1));
    if (have_initial)
        for (int32_t j = 0; j < 2 * n; ++j) kp2[j] = (float)(kp2[j] * kScales[3]); /* viso.cpp:367-370 */
    for (int i = 3; i >= 0; --i) {
        for (int32_t j = 0; j < 2 * n; ++j) kp1_[j] = (float)(kp1[j] * kScales[i]); /* viso.cpp:373-378 */
        lk_flow_single(c, pyr1 + i, pyr2 + i, n, kp1_, kp2, have_initial || i != 3, inverse, success);
        if (i != 0)
            for (int32_t j = 0; j < 2 * n; ++j) kp2[j] = (float)(kp2[j] / pyramid_scale); /* viso.cpp:384-389 */
    }
    free(kp1_);
}

/* include/bundle_adjuster.h:70-101  EdgeDirectProjection::computeError */
void photo_edge_errors(const direct_image* images, int32_t n, const double* points, const int32_t* src_slot,
                       const int32_t* tgt_slot, const double* frame_poses, const double* vertex_poses,
                       const double K[4], double* errors, uint8_t* level)
{
    const double fx = K[0], fy = K[1], cx = K[2], cy = K[3];
    for (int32_t e = 0; e < n; ++e) {
        const direct_image* src = images + 4 * (size_t)src_slot[e];
        const direct_image* tgt = images + 4 * (size_t)tgt_slot[e];
        const double* p = points + 3 * (size_t)e;
        const double* T = vertex_poses + 7 * (size_t)tgt_slot[e]; /* vertex(1), :73 */
        double R[9];
        oracle_quat_to_R(T, R); /* :74-75 */
        const double x = R[0] * p[0] + R[1] * p[1] + R[2] * p[2] + T[4];
        const double y = R[3] * p[0] + R[4] * p[1] + R[5] * p[2] + T[5];
        const double z = R[6] * p[0] + R[7] * p[1] + R[8] * p[2] + T[6];
        /* target_uv1 = K * (R p + t); target_uv1 /= z  (:77-78): K is the full 3x3 [fx 0 cx; 0 fy cy; 0 0 1] */
        const double tu = (fx * x + 0.0 * y + cx * z) / z, tv = (0.0 * x + fy * y + cy * z) / z;
        double suv[2];
        direct_cfg c;
        direct_default_cfg(&c);
        c.fx = fx, c.fy = fy, c.cx = cx, c.cy = cy;
        direct_project(&c, frame_poses + 7 * (size_t)src_slot[e], p, 1.0, suv); /* srcFrame->Project(p), :79 */
        double* err = errors + 16 * (size_t)e;
        for (int k = 0; k < 16; ++k) err[k] = 0.0;
        level[e] = 0;
        int idx = 0, abort_ = 0;
        for (int j = -2; (j <= 1) && !abort_; ++j) {
            for (int i = -2; i <= 1; ++i) {
                const float u1 = (float)(tu + (double)i), v1 = (float)(tv + (double)j);         /* :86-87 */
                const float u2 = (float)(suv[0] + (double)i), v2 = (float)(suv[1] + (double)j); /* :88-89 */
                if (!is_inside(tgt, u1, v1) || !is_inside(src, u2, v2)) { /* :91-95 */
                    level[e] = 1;
                    abort_ = 1;
                    break;
                }
                err[idx] = direct_pixel_value(src, u2, v2) - direct_pixel_value(tgt, u1, v1); /* :97 */
                ++idx;
            }
        }
    }
}
