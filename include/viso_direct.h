/*
 * viso_direct.h — C ABI of the B200-native direct (photometric) pose estimator of viso's tracker.
 *
 * the tracker's hot loop next to the bundle adjuster.  It replaces
 *   Viso::DirectPoseEstimationSingleLayer   reference src/viso.cpp:661-758
 *   Viso::DirectPoseEstimationMultiLayer    reference src/viso.cpp:760-766
 * and uses, with the reference's exact arithmetic,
 *   Keyframe::Project(point, level)         reference include/keyframe.h:82-89
 *   Keyframe::IsInside(u, v, level)         reference include/keyframe.h:76-79
 *   Keyframe::GetPixelValue (bilinear, u8)  reference include/keyframe.h:50-57
 *   Keyframe::GetGradient (central diff)    reference include/keyframe.h:59-64
 *   dPixeldXi (2x6, columns v then w)       reference src/viso.cpp:640-658
 *   T21 <- exp(update) * T21                reference src/viso.cpp:737
 *
 * The reference has no FFI: these are methods of class Viso that read Map, last_frame and the
 * frames' image pyramids.  The C ABI takes the same data as plain arrays; INTEGRATION.md shows
 * the call-site patch.  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions: pose = world->camera, 7 doubles (qx,qy,qz,qw, tx,ty,tz), Pc = R(q) Pw + t.
 * Images are 8-bit single channel (cv::Mat CV_8UC1 as Keyframe stores them), row stride `step`.
 *
 * There is NO CPU fallback: compute entry points return VISO_DIRECT_ERR_CUDA without a usable
 * sm_100 device.
 */
#ifndef VISO_DIRECT_H
#define VISO_DIRECT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VISO_DIRECT_MAX_LEVELS 4 /* Keyframe::nr_pyramids_, scales 1, 0.5, 0.25, 0.125 (viso.cpp:666) */

enum {
    VISO_DIRECT_OK = 0,
    VISO_DIRECT_ERR_CAPACITY = 1,
    VISO_DIRECT_ERR_CUDA = 3,
    VISO_DIRECT_ERR_NONFINITE = 5,
    VISO_DIRECT_ERR_ARG = 6,
    VISO_DIRECT_ERR_STATE = 7
};

typedef struct viso_direct viso_direct; /* opaque: device images, points, state, one stream */

typedef struct {
    int32_t struct_size;     /* = sizeof(viso_direct_config)                                     */
    int32_t device;          /* CUDA device ordinal                                              */
    int32_t max_points;      /* capacity of the map-point pool (Map::GetPoints(), map.h:15-19)   */
    int32_t max_frames;      /* how many (last_frame, current_frame, T21) problems run at once:
                                1 = the reference's call; > 1 = independent sequences batched     */
    int32_t max_width;       /* level-0 image size the device buffers are sized for              */
    int32_t max_height;
    double fx, fy, cx, cy;   /* shared K (viso.cpp:10); the reference passes K to dPixeldXi       */
    int32_t half_patch_size; /* Viso::half_patch_size = 4 (include/viso.h:25): 8x8 patch          */
    int32_t max_iterations;  /* 100 (viso.cpp:671)                                               */
    double delta_thresh;     /* 0.005 (viso.cpp:668)                                             */
    int32_t reference_termination;
                             /* 1 (default): the reference's loop exactly as written, viso.cpp:673-757 —
                                `cost` is never reset between iterations and `lastCost` starts at 0, so
                                (1 - cost/lastCost) is -inf after the first iteration and every level
                                stops after ONE Gauss-Newton step.
                                0: the loop the author evidently meant — cost reset per iteration, the
                                relative-decrease test skipped on iteration 0.                         */
    int32_t reserved;
} viso_direct_config;

/* what one level did (viso.cpp:673-757), per frame */
typedef struct {
    int32_t iterations;  /* Gauss-Newton updates applied to T21 on this level (before a restore) */
    int32_t n_good;      /* nGood of the last evaluated iteration (viso.cpp:717)                 */
    int32_t stop_reason; /* 0 = ran out of iterations, 1 = NaN update -> best restored (:741-744),
                            2 = cost went up -> best restored (:746-749), 3 = converged (:751-753) */
    int32_t reserved;
    double cost;         /* the value of `cost` when the loop ended                             */
    double last_cost;
} viso_direct_level_stats;

void viso_direct_default_config(viso_direct_config* cfg);
viso_direct* viso_direct_create(const viso_direct_config* cfg, int32_t* status);
void viso_direct_destroy(viso_direct* h);
const char* viso_direct_last_error(const viso_direct* h);
int64_t viso_direct_kernel_launches(const viso_direct* h);

/*
 * One pyramid level of one frame's image (Keyframe::Pyramids()[level], keyframe.h:29-45).
 * which: 0 = last_frame (the reference image), 1 = current_frame.  frame: 0 .. max_frames-1.
 * Reads by the reference that fall outside the cv::Mat buffer (undefined behaviour there: the
 * gradient taps of a patch that touches the border) read a clamped address here and in the oracle.
 */
int32_t viso_direct_set_image(viso_direct* h, int32_t frame, int32_t which, int32_t level,
                              const uint8_t* data, int32_t cols, int32_t rows, int32_t step);

/*
 * Keyframe's constructor (keyframe.h:29-45): `data` becomes level 0 and levels 1..3 are built on the device
 * with cv::pyrDown's arithmetic (bit-exact against OpenCV; see viso_lk_set_frame).
 */
int32_t viso_direct_set_frame(viso_direct* h, int32_t frame, int32_t which, const uint8_t* data,
                              int32_t cols, int32_t rows, int32_t step);

/* Map points in world coordinates (MapPoint::GetWorldPos, map_point.h:21-31), shared by all frames
 * of the handle when frame < 0, else the points of that frame's problem. */
int32_t viso_direct_set_points(viso_direct* h, int32_t frame, const double* xyz, int32_t n);

/* last_frame's pose (Keyframe::GetR/GetT), used by frame->Project(P1, level) at viso.cpp:693-695 */
int32_t viso_direct_set_reference_pose(viso_direct* h, int32_t frame, const double pose[7]);

/*
 * Viso::DirectPoseEstimationSingleLayer(level, current_frame, T21) for every frame of the handle.
 * poses: [n_frames][7], in = initial T21, out = refined T21.  stats (optional): [n_frames].
 */
int32_t viso_direct_estimate_level(viso_direct* h, int32_t level, int32_t n_frames, double* poses,
                                   viso_direct_level_stats* stats);

/*
 * Viso::DirectPoseEstimationMultiLayer: levels 3, 2, 1, 0 (viso.cpp:763-765).
 * stats (optional): [n_frames][VISO_DIRECT_MAX_LEVELS], indexed by level.
 * device_ms (optional): CUDA-event time of the kernels.
 */
int32_t viso_direct_estimate(viso_direct* h, int32_t n_frames, double* poses,
                             viso_direct_level_stats* stats, float* device_ms);

/*
 * The normal equations of ONE linearisation at the given poses (no update): H 6x6 row-major,
 * b 6, cost = sum of squared errors, n_good.  This is the loop body viso.cpp:688-733; exposed so
 * that parity tests can compare the accumulation itself with the oracle.
 */
int32_t viso_direct_linearize(viso_direct* h, int32_t level, int32_t n_frames, const double* poses,
                              double* H, double* b, double* cost, int32_t* n_good);

#ifdef __cplusplus
}
#endif
#endif
