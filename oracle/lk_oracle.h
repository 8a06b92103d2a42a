/*
 * lk_oracle.h — CPU restatement of viso's Lucas-Kanade patch alignment (TEST INFRASTRUCTURE ONLY).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load this.  Follows
 *   src/viso.cpp:845-925   Viso::LKAlignmentSingle            (inverse compositional, 100 iterations)
 *   src/viso.cpp:828-830   the level loop of Viso::LKAlignment (levels 3, 2, 1, 0)
 *   src/viso.cpp:259-351   Viso::OpticalFlowSingleLevel        (float keypoints, 10 iterations)
 *   src/viso.cpp:353-391   Viso::OpticalFlowMultiLevel
 *   include/common.h:35-49, include/keyframe.h:50-64   GetPixelValue, GetImageGradient / GetGradient
 * Parity unpinned: the reference has no tests for these and cannot be built here.
 */
#ifndef LK_ORACLE_H
#define LK_ORACLE_H

#include "direct_oracle.h" /* direct_image, direct_pixel_value, direct_gradient */

typedef struct {
    int32_t half_patch_size;         /* 4 (include/viso.h:25)                       */
    int32_t align_iterations;        /* 100 (src/viso.cpp:849)                      */
    int32_t flow_iterations;         /* 10 (src/viso.cpp:266)                       */
    int32_t reserved;
    double photometric_error_thresh; /* (2*4)*(2*4)*15*15 = 14400 (include/viso.h:26) */
} lk_cfg;

void lk_default_cfg(lk_cfg* c);

/* viso.cpp:845-925 for one level.  ref_images: [n_slots][4] pyramids; ref_slot[i] picks pair i's
 * reference keyframe; cur: the current frame's pyramid [4].  uv_cur is updated in place (:916). */
void lk_align_single(const lk_cfg* c, const direct_image* ref_images, const int32_t* ref_slot, const direct_image* cur,
                     int32_t n, const double* uv_ref, double* uv_cur, uint8_t* success, int32_t level);

/* viso.cpp:828-830 */
void lk_align(const lk_cfg* c, const direct_image* ref_images, const int32_t* ref_slot, const direct_image* cur,
              int32_t n, const double* uv_ref, double* uv_cur, uint8_t* success);

/* viso.cpp:259-351: kp1, kp2 = cv::KeyPoint::pt (float).  have_initial = !kp2.empty(). */
void lk_flow_single(const lk_cfg* c, const direct_image* img1, const direct_image* img2, int32_t n,
                    const float* kp1, float* kp2, int32_t have_initial, int32_t inverse, uint8_t* success);

/* viso.cpp:353-391: pyr1 / pyr2 = the two frames' pyramids [4] */
void lk_flow_multi(const lk_cfg* c, const direct_image* pyr1, const direct_image* pyr2, int32_t n,
                   const float* kp1, float* kp2, int32_t have_initial, int32_t inverse, uint8_t* success);

/* include/bundle_adjuster.h:70-101  EdgeDirectProjection::computeError for n edges.
 * images: [n_slots][4] pyramids (level 0 is used); poses [n_slots][7]; K = (fx, fy, cx, cy). */
void photo_edge_errors(const direct_image* images, int32_t n, const double* points, const int32_t* src_slot,
                       const int32_t* tgt_slot, const double* frame_poses, const double* vertex_poses,
                       const double K[4], double* errors, uint8_t* level);

#endif
