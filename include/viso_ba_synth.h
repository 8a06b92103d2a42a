/*
 * viso_ba_synth.h — deterministic synthetic bundle-adjustment problems.
 *
 * Host-only (no CUDA): built as libviso_synth.so from viso_b200/csrc/synth.cpp, so that the CPU
 * baseline / reference arm of bench.py and the CPU tests can generate the BASELINE.json
 * configurations without loading the CUDA backend.  Not part of the drop-in boundary
 * (include/viso_ba.h); the reference has no generator (it reads rgb/<id>.png, frame_sequence.h:25-38).
 */
#ifndef VISO_BA_SYNTH_H
#define VISO_BA_SYNTH_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int32_t struct_size;
    int32_t layout;        /* 0 = VO window on a line, 1 = 2-D camera grid, 2 = long trajectory */
    int64_t n_poses;
    int64_t n_points;
    int64_t target_edges;
    uint64_t seed;
    double fx, fy, cx, cy;
    double width, height, margin; /* 640 x 480 (main.cpp:73), 4 px (viso.h:25)              */
    double baseline;              /* 0.05, mean depth 1 (viso.cpp:622-637)                  */
    double noise_px;              /* 1.0                                                    */
    double outlier_frac;          /* 0 or 0.05                                              */
    double pose_rot_sigma;        /* 0.01 rad                                               */
    double pose_trans_sigma;      /* 0.02                                                   */
    double point_sigma;           /* 0.02                                                   */
    int32_t shuffle_edges;        /* 1: emit edges in random order (exercises the sort)     */
    int32_t reserved;
} viso_ba_synth_spec;

typedef struct viso_ba_synth viso_ba_synth;

/* config_id 1..5 = BASELINE.json configs; window = index inside config 5 */
int32_t viso_ba_synth_preset(int32_t config_id, int32_t window, viso_ba_synth_spec* out);
int32_t viso_ba_synth_create(const viso_ba_synth_spec* spec, viso_ba_synth** out);
void viso_ba_synth_destroy(viso_ba_synth* s);
int64_t viso_ba_synth_num_edges(const viso_ba_synth* s);
/* any output pointer may be NULL */
int32_t viso_ba_synth_fetch(const viso_ba_synth* s, double* poses_init7, double* poses_true7,
                            double* points_init3, double* points_true3, int32_t* edge_pose,
                            int32_t* edge_point, float* edge_uv);

#ifdef __cplusplus
}
#endif
#endif /* VISO_BA_SYNTH_H */
