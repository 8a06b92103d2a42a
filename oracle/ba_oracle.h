/*
 * ba_oracle.h — CPU oracle for the viso bundle-adjustment hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under viso_b200/ may include, link or
 * call this.  Allowed callers: tests/, __graft_entry__.smoke(), and the
 * cpu_baseline / --impl reference legs of bench.py.
 *
 * PARITY UNPINNED: the reference's own adjuster (include/bundle_adjuster.h)
 * does not compile (undeclared identifiers at :157-159), is included by no
 * translation unit, has empty AddPoint/AddPose/AddEdge bodies (:169-173) and
 * delegates all arithmetic to g2o + Sophus, neither of which is vendored,
 * pinned or installed here.  The reference has no tests or golden vectors.
 * This file is therefore a plain-C fp64 restatement of
 *   - the reference's conventions where it states them (cited per function),
 *   - g2o's published Levenberg-Marquardt / BlockSolver<6,3> / Huber
 *     algorithm (version unpinned; ~June-2018 master) where it does not.
 */
#ifndef VISO_BA_ORACLE_H
#define VISO_BA_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    double fx, fy, cx, cy; /* shared pinhole K: src/main.cpp:14-17, src/viso.cpp:10 */
    double huber_delta;    /* <= 0 : plain L2 (reference sets no robust kernel, bundle_adjuster.h:160) */
    double tau;            /* g2o OptimizationAlgorithmLevenberg _tau, default 1e-5 */
    double lambda_init;    /* > 0 : user lambda; else tau * max diag(H) */
    int32_t max_trials;    /* g2o _maxTrialsAfterFailure, default 10 */
    int32_t threads;       /* OpenMP threads; 1 = the reference's single-threaded shape */
} oracle_cfg;

typedef struct {
    double cost_before; /* robustified chi2 at the start of the iteration   */
    double cost_after;  /* after the accepted step (== before if none)      */
    double lambda;      /* damping after the iteration                      */
    int32_t trials;     /* solve+evaluate trials spent in this iteration    */
    int32_t accepted;   /* 1 if a step was accepted                         */
} oracle_iter_stats;

void oracle_default_cfg(oracle_cfg* c);

/* R (row-major 3x3) from unit quaternion (x,y,z,w). */
void oracle_quat_to_R(const double q[4], double R[9]);

/* keyframe.h:82-89  Pc = R*Pw + T ; (u,v) = (fx*x/z + cx, fy*y/z + cy), level 0. */
void oracle_project(const oracle_cfg* c, const double pose[7], const double Pw[3],
                    double uv[2], double Pc[3]);

/* viso.cpp:640-658 dPixeldXi, scale = 1.  J is row-major 2x6, columns (v, w). */
void oracle_jac_pose(const oracle_cfg* c, const double Pc[3], double J[12]);

/* first three columns of dPixeldXi times R.  J row-major 2x3. */
void oracle_jac_point(const oracle_cfg* c, const double Pc[3], const double R[9], double J[6]);

/* Sophus SE3d::exp: xi = (upsilon, omega) -> (q xyzw, t). */
void oracle_se3_exp(const double xi[6], double pose[7]);

/* bundle_adjuster.h:51-55  T <- exp(xi) * T (left-multiplicative). */
void oracle_pose_oplus(const double xi[6], double pose[7]);

/* g2o RobustKernelHuber: returns rho(s), writes w = rho'(s); s = |r|^2. */
double oracle_huber(double s, double delta, double* w);

/* robustified chi2 over all observations (g2o activeRobustChi2). */
double oracle_cost(const oracle_cfg* c, int32_t n_poses, const double* poses,
                   int32_t n_points, const double* points, int64_t n_obs,
                   const int32_t* obs_pose, const int32_t* obs_point, const float* obs_uv);

/*
 * One linearisation at the given state and damping: the reduced camera system
 * S (row-major n x n, n = 6 * #free poses), its right-hand side bs (n) and the
 * maximum diagonal entry of the undamped H over free vertices.  free_index
 * (n_poses ints, caller-provided) receives pose -> free slot or -1.
 * Returns n, or a negative value on bad input.
 */
int32_t oracle_reduced_system(const oracle_cfg* c, int32_t n_poses, const double* poses,
                              const uint8_t* pose_fixed, int32_t n_points, const double* points,
                              int64_t n_obs, const int32_t* obs_pose, const int32_t* obs_point,
                              const float* obs_uv, double lambda, double* S, double* bs,
                              double* max_diag, int32_t* free_index);

/*
 * Levenberg-Marquardt (g2o policy) for `iterations` outer iterations.
 * poses (7 per pose) and points (3 per point) are updated in place.
 * stats may be NULL; otherwise it has room for `iterations` entries.
 * Returns the number of outer iterations executed (>= 0) or < 0 on error.
 */
int32_t oracle_optimize(const oracle_cfg* c, int32_t n_poses, double* poses,
                        const uint8_t* pose_fixed, int32_t n_points, double* points,
                        int64_t n_obs, const int32_t* obs_pose, const int32_t* obs_point,
                        const float* obs_uv, int32_t iterations, oracle_iter_stats* stats,
                        double* final_cost);

/* dense in-place Cholesky (lower, row-major) + solve; returns 0 ok, 1 not PD. */
int32_t oracle_cholesky_solve(int32_t n, double* A, double* b, int32_t threads);

#ifdef __cplusplus
}
#endif
#endif
