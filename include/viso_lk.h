/*
 * viso_lk.h — C ABI of the B200-native Lucas-Kanade patch alignment of viso's front end.
 *
 * SURVEY.md section 8 row f4.  It replaces
 *   Viso::LKAlignmentSingle        reference src/viso.cpp:845-925   (2x2 Gauss-Newton on an 8x8 patch,
 *                                                                     inverse compositional, per pair)
 *   the level loop of LKAlignment  reference src/viso.cpp:828-830   (levels 3, 2, 1, 0)
 *   Viso::OpticalFlowSingleLevel   reference src/viso.cpp:259-351
 *   Viso::OpticalFlowMultiLevel    reference src/viso.cpp:353-391
 * with the reference's samplers GetPixelValue / GetImageGradient (include/common.h:35-49,
 * include/keyframe.h:50-64).  Every keypoint is independent: one warp per keypoint.
 *
 * The reference has no FFI (methods of class Viso reading cv::Mat pyramids); the C ABI takes the
 * same data as plain arrays.  No C++/torch types.  NO CPU fallback.
 */
#ifndef VISO_LK_H
#define VISO_LK_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VISO_LK_LEVELS 4 /* nr_pyramids (viso.cpp:362), scales 1, 0.5, 0.25, 0.125 (viso.cpp:364, :850) */

enum {
    VISO_LK_OK = 0,
    VISO_LK_ERR_CAPACITY = 1,
    VISO_LK_ERR_INDEX = 2,
    VISO_LK_ERR_CUDA = 3,
    VISO_LK_ERR_NONFINITE = 5,
    VISO_LK_ERR_ARG = 6,
    VISO_LK_ERR_STATE = 7
};

typedef struct viso_lk viso_lk; /* opaque: device pyramids of the image slots, keypoint buffers, one stream */

typedef struct {
    int32_t struct_size;             /* = sizeof(viso_lk_config)                                   */
    int32_t device;
    int32_t max_keypoints;           /* pairs / keypoints per call                                 */
    int32_t max_slots;               /* image slots: the map's keyframes + the current frame       */
    int32_t max_width, max_height;   /* level-0 image size                                         */
    int32_t half_patch_size;         /* 4 (include/viso.h:25)                                      */
    int32_t align_iterations;        /* 100 (viso.cpp:849)                                         */
    int32_t flow_iterations;         /* 10 (viso.cpp:266)                                          */
    int32_t reserved;
    double photometric_error_thresh; /* (2*4)*(2*4)*15*15 = 14400 (include/viso.h:26)              */
} viso_lk_config;

void viso_lk_default_config(viso_lk_config* cfg);
viso_lk* viso_lk_create(const viso_lk_config* cfg, int32_t* status);
void viso_lk_destroy(viso_lk* h);
const char* viso_lk_last_error(const viso_lk* h);
int64_t viso_lk_kernel_launches(const viso_lk* h);

/* One pyramid level of the image in `slot` (Keyframe::Pyramids()[level], keyframe.h:29-45). */
int32_t viso_lk_set_image(viso_lk* h, int32_t slot, int32_t level, const uint8_t* data,
                          int32_t cols, int32_t rows, int32_t step);

/*
 * Keyframe's constructor (keyframe.h:29-45): `data` becomes level 0 of `slot` and levels 1..3 are built on
 * the device with cv::pyrDown's arithmetic (5-tap [1 4 6 4 1], BORDER_REFLECT_101, (sum + 128) >> 8,
 * Size(cols * 0.5, rows * 0.5)), bit-exact against OpenCV.
 */
int32_t viso_lk_set_frame(viso_lk* h, int32_t slot, const uint8_t* data, int32_t cols, int32_t rows, int32_t step);

/* Keyframe::Pyramids()[level] of `slot` back on the host, rows packed (stride = cols); capacity in bytes. */
int32_t viso_lk_get_image(viso_lk* h, int32_t slot, int32_t level, uint8_t* out, int64_t capacity,
                          int32_t* cols, int32_t* rows);

/*
 * The level loop of Viso::LKAlignment (viso.cpp:828-830) over LKAlignmentSingle (viso.cpp:845-925).
 * Pair i = AlignmentPair{ref_frame = slot ref_slot[i], cur_frame = slot cur_slot, uv_ref, uv_cur}
 * (include/viso.h:116-121).  uv_ref [n][2] level-0 pixels; uv_cur [n][2] in = projection of the map
 * point into the current frame (viso.cpp:816), out = aligned position (viso.cpp:916); success [n] =
 * the flags of the finest level, which is what LKAlignment keeps (viso.cpp:832-843).
 */
int32_t viso_lk_align(viso_lk* h, int32_t n, const int32_t* ref_slot, int32_t cur_slot,
                      const double* uv_ref, double* uv_cur, uint8_t* success, float* device_ms);

/* LKAlignmentSingle on ONE level (viso.cpp:845-925) */
int32_t viso_lk_align_level(viso_lk* h, int32_t level, int32_t n, const int32_t* ref_slot, int32_t cur_slot,
                            const double* uv_ref, double* uv_cur, uint8_t* success);

/*
 * Viso::OpticalFlowMultiLevel(ref_frame = slot1, cur_frame = slot2, kp1, kp2, success, inverse)
 * (viso.cpp:353-391).  kp1, kp2: cv::KeyPoint::pt as [n][2] float; have_initial = !kp2.empty().
 */
int32_t viso_lk_optical_flow(viso_lk* h, int32_t slot1, int32_t slot2, int32_t n, const float* kp1,
                             float* kp2, int32_t have_initial, int32_t inverse, uint8_t* success,
                             float* device_ms);

/* Viso::OpticalFlowSingleLevel(img1 = slot1 level, img2 = slot2 level, ...) (viso.cpp:259-351) */
int32_t viso_lk_optical_flow_level(viso_lk* h, int32_t level, int32_t slot1, int32_t slot2, int32_t n,
                                   const float* kp1, float* kp2, int32_t have_initial, int32_t inverse,
                                   uint8_t* success);

/*
 * EdgeDirectProjection::computeError (reference include/bundle_adjuster.h:70-101), the reference adjuster's
 * only edge, for n edges at once: the 16-D photometric residual over the 4x4
 * patch (i, j in [-2, 1]) between srcFrame (slot src_slot[e], projected with that keyframe's own pose
 * frame_poses[src_slot[e]], :79) and targetFrame (slot tgt_slot[e], projected with the pose VERTEX
 * vertex_poses[tgt_slot[e]] through K, :74-78), pixel coordinates cast to float as at :86-89.
 * errors [n][16] in the reference's idx order (j outer, i inner, :83-99); level[e] = 1 when a sample
 * fell outside an image — the edge disables itself with setLevel(1) and stops (:91-95); its remaining
 * entries are then left at 0 here (the reference leaves stale values).  Poses [max_slots][7].
 */
int32_t viso_lk_edge_direct_projection(viso_lk* h, int32_t n, const double* points, const int32_t* src_slot,
                                       const int32_t* tgt_slot, const double* frame_poses,
                                       const double* vertex_poses, double fx, double fy, double cx, double cy,
                                       double* errors, uint8_t* level, float* device_ms);

#ifdef __cplusplus
}
#endif
#endif
