/*
 * direct_oracle.h — CPU restatement of viso's direct pose estimation (TEST INFRASTRUCTURE ONLY).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load this.  It follows the
 * reference statement by statement (file:line cited at every function in direct_oracle.c):
 *   src/viso.cpp:661-758   Viso::DirectPoseEstimationSingleLayer
 *   src/viso.cpp:760-766   Viso::DirectPoseEstimationMultiLayer
 *   src/viso.cpp:640-658   dPixeldXi
 *   include/keyframe.h:50-64, :76-89   GetPixelValue, GetGradient, IsInside, Project
 * Parity unpinned: the reference has no tests or fixtures for this function and cannot be built
 * here (OpenCV, Eigen, Sophus absent).  The restatement is cross-checked by an independent numpy
 * statement (tests/numpy_direct.py) and by recovering a known camera motion on synthetic images.
 */
#ifndef DIRECT_ORACLE_H
#define DIRECT_ORACLE_H

#include <stdint.h>

typedef struct {
    const uint8_t* data; /* cv::Mat::data of pyramids_[level] */
    int32_t cols, rows, step;
} direct_image;

typedef struct {
    double fx, fy, cx, cy;
    int32_t half_patch_size;       /* 4  (include/viso.h:25)  */
    int32_t max_iterations;        /* 100 (src/viso.cpp:671)  */
    double delta_thresh;           /* 0.005 (src/viso.cpp:668) */
    int32_t reference_termination; /* 1: the loop exactly as written; 0: cost reset per iteration */
    int32_t reserved;
} direct_cfg;

typedef struct {
    int32_t iterations, n_good, stop_reason, reserved;
    double cost, last_cost;
} direct_level_stats;

void direct_default_cfg(direct_cfg* c);

/* keyframe.h:50-57 */
double direct_pixel_value(const direct_image* im, double x, double y);
/* keyframe.h:59-64 */
void direct_gradient(const direct_image* im, double u, double v, double g[2]);
/* keyframe.h:82-89 with scales_[level] */
void direct_project(const direct_cfg* c, const double pose[7], const double Pw[3], double scale, double uv[2]);
/* viso.cpp:640-658 */
void direct_dpixel_dxi(const direct_cfg* c, const double R[9], const double T[3], const double P[3],
                       double scale, double J[12]);

/* loop body viso.cpp:688-733: H (6x6 row-major), b (6), sum of squared errors, nGood */
void direct_linearize(const direct_cfg* c, int32_t level, const direct_image* ref, const direct_image* cur,
                      const double ref_pose[7], const double T21[7], const double* points, int32_t n_points,
                      double H[36], double b[6], double* cost, int32_t* n_good);

/* viso.cpp:661-758 */
void direct_single_layer(const direct_cfg* c, int32_t level, const direct_image* ref, const direct_image* cur,
                         const double ref_pose[7], double T21[7], const double* points, int32_t n_points,
                         direct_level_stats* stats);

/* viso.cpp:760-766; ref / cur: arrays of 4 levels; stats (optional): [4] indexed by level */
void direct_multi_layer(const direct_cfg* c, const direct_image* ref, const direct_image* cur,
                        const double ref_pose[7], double T21[7], const double* points, int32_t n_points,
                        direct_level_stats* stats);

#endif
