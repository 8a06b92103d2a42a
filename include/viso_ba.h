/*
 * viso_ba.h — C ABI of the B200-native bundle-adjustment backend for viso.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.
 * The reference has no FFI; its boundary is the header-only C++ class
 * `BundleAdjuster` (reference include/bundle_adjuster.h:23-187).  Each entry
 * point below cites the reference interface it replaces or completes; the
 * source-compatible C++ wrapper is include/bundle_adjuster.h in this repo.
 *
 * Conventions (all from the reference):
 *   pose   = world->camera, 7 doubles (qx,qy,qz,qw, tx,ty,tz)  [Sophus SE3d::data() order]
 *            Pc = R(q) * Pw + t                                 keyframe.h:82-89
 *   point  = 3 doubles, world frame                             bundle_adjuster.h:24-39
 *   edge   = (point, pose, u, v); residual r = project - (u,v)  bundle_adjuster.h:72-73, :158-159
 *   update = pose: T <- exp(d) T, d = (v, w)                    bundle_adjuster.h:51-55
 *            point: P <- P + d                                  bundle_adjuster.h:34-38
 *
 * There is NO CPU fallback: every compute entry point returns
 * VISO_BA_ERR_CUDA when no sm_100 device / driver is usable.
 */
#ifndef VISO_BA_H
#define VISO_BA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VISO_BA_VERSION 100

/* status codes (the reference prints to std::cerr and exit(0)s, bundle_adjuster.h:133-135,
 * :162-165; this ABI never exits and is quiet unless config.verbose) */
enum {
    VISO_BA_OK = 0,
    VISO_BA_ERR_CAPACITY = 1,  /* more points/poses/edges than the handle's pools    */
    VISO_BA_ERR_INDEX = 2,     /* edge refers to a point/pose that does not exist    */
    VISO_BA_ERR_CUDA = 3,      /* CUDA runtime / driver / device error               */
    VISO_BA_ERR_NCCL = 4,      /* NCCL missing or a collective failed                */
    VISO_BA_ERR_NONFINITE = 5, /* NaN/Inf in poses, points or measurements           */
    VISO_BA_ERR_ARG = 6,       /* NULL pointer, bad size, bad struct_size            */
    VISO_BA_ERR_STATE = 7,     /* call not valid in the handle's current state       */
    VISO_BA_ERR_UNSUPPORTED = 8
};

typedef struct viso_ba viso_ba; /* opaque; owns host staging + device memory + one stream */

/*
 * Replaces the compile-time pools of bundle_adjuster.h:182-184 (1000 each) and the
 * hard-coded intrinsics of src/main.cpp:14-17.  Fill with viso_ba_default_config().
 */
typedef struct {
    int32_t struct_size;   /* = sizeof(viso_ba_config)                                  */
    int32_t device;        /* CUDA device ordinal                                       */
    int64_t max_points;    /* pool capacities; default 1000 (bundle_adjuster.h:182-184) */
    int64_t max_poses;
    int64_t max_edges;
    double fx, fy, cx, cy; /* one shared K, as Viso sets on every frame (viso.cpp:10)    */
    double huber_delta;    /* <= 0: plain L2 (reference sets no robust kernel, :160)    */
    double lm_tau;         /* lambda_0 = tau * max diag(H); g2o default 1e-5            */
    double lm_lambda_init; /* > 0 overrides tau rule                                    */
    int32_t lm_max_trials; /* g2o maxTrialsAfterFailure, default 10                     */
    int32_t verbose;       /* reference: optimizer_.setVerbose(true) (:120); default 0  */
    int32_t fix_first_pose;/* gauge: fix the first pose of every problem (viso keeps frame 0
                              at identity, keyframe.h:34-35); default 1                  */
    int32_t use_cuda_graph;/* replay one captured LM trial instead of re-launching; default 1 */
    int32_t path;          /* which kernels linearise: 0 = auto (default), 1 = window fast path
                              whenever eligible (every problem <= 32 keyframes and <= 27 free
                              poses, no duplicate edges, not sharded), 2 = general pair-list
                              path.  Auto takes the window path whenever the problem is eligible. */
    int32_t solver_segments; /* reduced systems beyond 27 free poses (one large problem): how the pose chain is cut for the
                              solve.  0 = auto (up to 8 interior segments when the co-visibility span is small against
                              the chain, as on a long trajectory), 1 = never (one envelope Cholesky), k >= 2 = up to k
                              segments whenever the structure allows it.  The result does not depend on the GPU count. */
    int32_t host_structure_build; /* general path, one large problem: 0 (default) = the structure (sort by point and pose, pair
                              lists by S block) is built on the device from 200 k edges on, 1 = always on the host threads.
                              Both build the same tables. */
} viso_ba_config;

/* per outer LM iteration (single-problem handles) */
typedef struct {
    double cost_before;
    double cost_after;
    double lambda;
    int32_t trials;
    int32_t accepted;
} viso_ba_iter_stats;

typedef struct {
    double initial_cost;
    double final_cost;
    double final_lambda;
    int32_t iterations;   /* outer iterations executed                                  */
    int32_t trials;       /* linearise+solve+evaluate trials executed                   */
    int32_t terminated;   /* 1 if LM stopped early (max trials / rho == 0 / lambda inf) */
    int32_t reserved;
    float device_ms;      /* CUDA-event time of the LM loop on the handle's stream      */
    float setup_ms;       /* host+device time of upload / sort / structure build        */
} viso_ba_summary;

int32_t viso_ba_version(void);
const char* viso_ba_status_string(int32_t status);
void viso_ba_default_config(viso_ba_config* cfg);

/* BundleAdjuster::BundleAdjuster()  (bundle_adjuster.h:111-167): allocate the pools once */
int32_t viso_ba_create(const viso_ba_config* cfg, viso_ba** out);
void viso_ba_destroy(viso_ba* h);
const char* viso_ba_last_error(const viso_ba* h);

/* drop every point / pose / edge, keep pools and config */
int32_t viso_ba_clear(viso_ba* h);

/* int AddPoint(V3d)            (bundle_adjuster.h:169): returns the new point index, or -status */
int64_t viso_ba_add_point(viso_ba* h, const double xyz[3]);
/* int AddPose(Sophus::SE3d)    (bundle_adjuster.h:171): returns the new pose index, or -status  */
int64_t viso_ba_add_pose(viso_ba* h, const double q_xyzw_t[7]);
/* void AddEdge(int point,int pose) (bundle_adjuster.h:173), completed with the measured pixel
 * (Keypoints()[idx].pt is a cv::Point2f, keyframe.h:15 / map_point.h:21-24 -> float32) */
int32_t viso_ba_add_edge(viso_ba* h, int64_t point, int64_t pose, float u, float v);

/* bulk form of the three calls above: replaces the whole problem in one copy */
int32_t viso_ba_set_problem(viso_ba* h, int64_t n_poses, const double* poses7,
                            int64_t n_points, const double* points3, int64_t n_edges,
                            const int32_t* edge_pose, const int32_t* edge_point,
                            const float* edge_uv);

/*
 * batched independent windows (BASELINE config 5): problem b owns poses
 * [pose_off[b], pose_off[b+1]), points [point_off[b], ..) and edges [edge_off[b], ..);
 * edge_pose / edge_point are indices LOCAL to the problem.  Each problem runs its own LM.
 */
int32_t viso_ba_set_problem_batched(viso_ba* h, int64_t n_problems, const int64_t* pose_off,
                                    const int64_t* point_off, const int64_t* edge_off,
                                    const double* poses7, const double* points3,
                                    const int32_t* edge_pose, const int32_t* edge_point,
                                    const float* edge_uv);

/*
 * The reference adjuster's OWN edge: EdgeDirectProjection(srcFrame, targetFrame) between VertexPoint and VertexPose
 * (bundle_adjuster.h:58-106, :154-160): the 16-D photometric residual of a 4 x 4 patch, information I_16.
 *   - every pose vertex has a keyframe = its level-0 image (CV_8UC1, keyframe.h:29-45): viso_ba_set_pose_image
 *   - an edge names a point, a SOURCE keyframe (by pose index) and the pose vertex / target keyframe; the source
 *     projects with the pose its keyframe had at upload (srcFrame->Project, :79), the target with the vertex (:73-78)
 *   - an edge whose patch leaves an image disables itself at that state (setLevel(1), :91-95)
 *   - the Jacobian, which the reference leaves to g2o's numeric differentiation, is analytic: image gradient
 *     (keyframe.h:59-64) x dPixeldXi (viso.cpp:640-658) for the pose, x the projection Jacobians of both frames for the point
 * A handle holds either reprojection or photometric edges (VISO_BA_ERR_UNSUPPORTED otherwise); photometric
 * problems are single, unsharded, and run on the general kernels (config.path is ignored).
 * Photometric edges leave the scale of a window free: fix two poses (viso_ba_set_pose_fixed) to pin it.
 */
int32_t viso_ba_set_pose_image(viso_ba* h, int64_t pose, const uint8_t* data, int32_t rows, int32_t cols,
                               int32_t stride);
int32_t viso_ba_add_photo_edge(viso_ba* h, int64_t point, int64_t src_pose, int64_t dst_pose);
/* bulk form: replaces the whole problem (images are set afterwards, per pose) */
int32_t viso_ba_set_photo_problem(viso_ba* h, int64_t n_poses, const double* poses7, int64_t n_points,
                                  const double* points3, int64_t n_edges, const int32_t* edge_point,
                                  const int32_t* edge_src, const int32_t* edge_dst);

/* gauge: g2o setFixed(); the reference fixes nothing, viso keeps frame 0 at identity
 * (keyframe.h:34-35, viso.cpp:82-83).  Batched handles: global pose index. */
int32_t viso_ba_set_pose_fixed(viso_ba* h, int64_t pose, int32_t fixed);
int32_t viso_ba_set_intrinsics(viso_ba* h, double fx, double fy, double cx, double cy);
int32_t viso_ba_set_huber_delta(viso_ba* h, double delta);

/*
 * Shard one problem across the GPUs of one box: every rank passes the SAME full problem;
 * rank r keeps a contiguous range of points (balanced on edge count) and all poses.  The
 * reduced camera system is combined with one fp64 ncclAllReduce per linearisation.
 * nccl_id: 128 bytes from viso_ba_nccl_unique_id() on rank 0, broadcast by the caller.
 */
int32_t viso_ba_nccl_unique_id(void* out128);
int32_t viso_ba_shard_init(viso_ba* h, int32_t rank, int32_t n_ranks, const void* nccl_id128);

/* the partition viso_ba_shard_init uses (host-only, needs no GPU): rank r owns the points
 * [out_bounds[r], out_bounds[r+1]); contiguous ranges balanced on edge count */
int32_t viso_ba_shard_point_ranges(int64_t n_points, int64_t n_edges, const int32_t* edge_point,
                                   int32_t n_ranks, int64_t* out_bounds);

/* run on a caller-provided cudaStream_t (NULL: the handle's own stream) */
int32_t viso_ba_set_stream(viso_ba* h, void* cuda_stream);

/*
 * The step the reference never reached (optimizer_.initializeOptimization()/optimize()).
 *   upload   : H2D of the staged problem + sort by (point,pose) + block structure; idempotent
 *   optimize : upload if needed, then `iterations` LM iterations on the device, then D2H
 *   reset    : restore the device state to the uploaded initial poses/points (bench loops)
 * stats may be NULL (else `iterations` entries, single-problem handles only).
 */
int32_t viso_ba_upload(viso_ba* h);
int32_t viso_ba_optimize(viso_ba* h, int32_t iterations, viso_ba_iter_stats* stats,
                         viso_ba_summary* summary);
int32_t viso_ba_reset_state(viso_ba* h);
/* device-only LM loop: no host<->device copies of poses/points (bench `value` leg) */
int32_t viso_ba_optimize_resident(viso_ba* h, int32_t iterations, viso_ba_summary* summary);
/*
 * device state -> the handle's host copy: the read-back viso_ba_optimize ends with, as a call of its own.
 * upload / optimize_resident / download are viso_ba_optimize in three steps, for a caller that feeds a stream of
 * problems through two handles and wants only the LM loops to take turns on the GPU (viso_b200/stream.py).
 */
int32_t viso_ba_download(viso_ba* h);

/* read-back the reference never had */
int64_t viso_ba_num_points(const viso_ba* h);
int64_t viso_ba_num_poses(const viso_ba* h);
int64_t viso_ba_num_edges(const viso_ba* h);
int64_t viso_ba_num_problems(const viso_ba* h);
int32_t viso_ba_get_poses(viso_ba* h, int64_t first, int64_t count, double* out7);
int32_t viso_ba_get_points(viso_ba* h, int64_t first, int64_t count, double* out3);
/* batched handles: per-problem summaries (n_problems entries) */
int32_t viso_ba_get_problem_summaries(viso_ba* h, viso_ba_summary* out);

/* robustified cost of the current device state (one kernel pass; no LM) */
int32_t viso_ba_eval_cost(viso_ba* h, double* cost);

/*
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * The same robustified reprojection cost as viso_ba_eval_cost, of a raw edge list that is ALREADY on the device, computed
 * with no handle and no allocation: no cudaMalloc / cudaFree / allocating library call, here or inside.  The scratch the
 * pass needs (the R|t table of the poses, the partial sums of the CTAs) lives in one file-scope `static __device__`
 * array of 402,688 bytes (viso_b200/csrc/ba_static_cost.cu) that every GPU loading the library holds until exit.
 *   - runs on the calling thread's CURRENT device, on cuda_stream (a cudaStream_t of that device; NULL: the default
 *     stream); three kernels are enqueued and the call returns without synchronising: *cost_out_dev is valid once the
 *     stream has reached that point
 *   - every *_dev pointer is device memory of that device: poses7 [n_poses][7] (q xyzw, t), points3 [n_points][3],
 *     edge_pose / edge_point [n_edges], edge_uv [n_edges][2] float32 (the layout of viso_ba_set_problem), one double out
 *   - limits: n_poses <= VISO_BA_STATIC_MAX_POSES, n_edges <= VISO_BA_STATIC_MAX_EDGES; beyond them the call returns
 *     VISO_BA_ERR_CAPACITY, enqueues nothing and says so in err (err_cap bytes, may be NULL)
 *   - edge indices are checked on the device: one out of range makes the cost NaN (nothing is read out of bounds)
 *   - the sum has a fixed shape: repeated calls on the same input return the same bits; viso_ba_eval_cost sums in another
 *     order, so the two agree to rounding (~1e-15 relative)
 *   - the first call on a device looks the array's address up (cudaGetSymbolAddress): make that call outside a capture
 *   - every call on a GPU shares the array, which keeps the last call's contents (nothing in it needs clearing: a call
 *     writes each word before reading it).  Calls that OVERLAP on one GPU (two streams, two threads) race: serialise them.
 */
#define VISO_BA_STATIC_MAX_POSES 4096
#define VISO_BA_STATIC_MAX_EDGES 2147483647LL
int32_t viso_ba_eval_cost_static(void* cuda_stream, int64_t n_poses, const double* poses7_dev, int64_t n_points,
                                 const double* points3_dev, int64_t n_edges, const int32_t* edge_pose_dev,
                                 const int32_t* edge_point_dev, const float* edge_uv_dev, double fx, double fy,
                                 double cx, double cy, double huber_delta, double* cost_out_dev, char* err,
                                 int32_t err_cap);

/*
 * debug / parity hooks: one linearisation of the current state with damping lambda.
 * S is the dense reduced camera system (row-major n x n, n = 6 * #free poses of problem 0),
 * bs its right-hand side (b = -gradient convention of g2o).  Single-problem handles.
 */
int32_t viso_ba_debug_reduced_system(viso_ba* h, double lambda, int32_t n_cap, double* S,
                                     double* bs, int32_t* n_out);

/* debug / test hook: the host part of the general structure build (sort by point and pose, pair lists by S block, work
 * items) with `threads` host threads (<= 0: the default policy); needs no GPU.  The checksum covers every table it builds:
 * it must not depend on the thread count. */
int32_t viso_ba_debug_host_build(viso_ba* h, int32_t threads, double* ms_out, uint64_t* checksum_out);

/* launch accounting for bench.py: kernels launched by this handle since creation */
int64_t viso_ba_kernel_launches(const viso_ba* h);
/* algorithmic bytes / per-kernel device time of the last optimize call (bench roofline) */
typedef struct {
    char name[32];
    int64_t launches;
    double device_ms; /* CUDA-event total over the launches (only when profiling is on) */
} viso_ba_kernel_time;
int32_t viso_ba_set_kernel_timing(viso_ba* h, int32_t on);
int32_t viso_ba_get_kernel_times(viso_ba* h, int32_t cap, viso_ba_kernel_time* out,
                                 int32_t* n_out);

/* The synthetic problem generator is host-only test/bench infrastructure and lives in its
 * own library, libviso_synth.so: include/viso_ba_synth.h.  The CPU reference arm of bench.py uses it without
 * ever mapping this CUDA library. */

#ifdef __cplusplus
}
#endif
#endif /* VISO_BA_H */
