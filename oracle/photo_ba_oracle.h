/*
 * photo_ba_oracle.h — CPU oracle of bundle adjustment over EdgeDirectProjection, the photometric edge of the
 * reference adjuster (reference include/bundle_adjuster.h:58-106).  TEST INFRASTRUCTURE ONLY; see photo_ba_oracle.c.
 */
#ifndef PHOTO_BA_ORACLE_H
#define PHOTO_BA_ORACLE_H

#include "ba_oracle.h"
#include "direct_oracle.h"

#ifdef __cplusplus
extern "C" {
#endif

/* images: one level-0 image per pose (the keyframe of that pose vertex).  Edge e = (point e_point[e], source keyframe
 * e_src[e] with its constant pose src_poses[e_src[e]], pose vertex / target keyframe e_dst[e]). */
double photo_ba_cost(const oracle_cfg* c, const direct_image* images, int32_t n_poses, const double* poses,
                     const double* src_poses, const double* points, int64_t n_edges, const int32_t* e_point,
                     const int32_t* e_src, const int32_t* e_dst);

/* residual r (16) and the analytic Jacobians Jp (16 x 6, pose vertex) and Jl (16 x 3, point) of one edge; returns 0
 * when a sample leaves an image (the edge disables itself, bundle_adjuster.h:91-95) */
int32_t photo_ba_edge_jacobian(const oracle_cfg* c, const direct_image* src, const direct_image* tgt, const double src_pose[7],
                               const double dst_pose[7], const double p[3], double r[16], double Jp[96], double Jl[48]);

/* g2o Levenberg-Marquardt over the photometric edges; the source poses are frozen at the incoming `poses` */
int32_t photo_ba_optimize(const oracle_cfg* c, const direct_image* images, int32_t n_poses, double* poses,
                          const uint8_t* pose_fixed, int32_t n_points, double* points, int64_t n_edges,
                          const int32_t* e_point, const int32_t* e_src, const int32_t* e_dst, int32_t iterations,
                          oracle_iter_stats* stats, double* final_cost);

#ifdef __cplusplus
}
#endif
#endif
