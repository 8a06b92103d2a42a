// direct_pose.cu — B200-native direct (photometric) pose estimation of viso's tracker
// (SURVEY.md section 8 row f3).  C ABI: include/viso_direct.h.
//
// Reference semantics (citations relative to /root/reference):
//   Viso::DirectPoseEstimationSingleLayer  src/viso.cpp:661-758   Gauss-Newton on one pyramid level
//   Viso::DirectPoseEstimationMultiLayer   src/viso.cpp:760-766   levels 3, 2, 1, 0
//   dPixeldXi                              src/viso.cpp:640-658   2x6, columns (v, w)
//   Keyframe::GetPixelValue / GetGradient  include/keyframe.h:50-64
//   Keyframe::IsInside / Project           include/keyframe.h:76-89
//
// B200 design: the whole multi-level Gauss-Newton loop of one frame is ONE kernel launch by ONE
// thread-block cluster (16 CTAs x 512 threads; 8 CTAs when many frames are batched).  A warp owns a point at a time and its 32 lanes own
// two pixels of the 8x8 patch each; every lane keeps the 21 + 6 + 1 sums (upper triangle of H, b,
// cost) in registers over all points of its warp.  The sums are folded with a fixed butterfly
// inside the warp, in warp order inside the CTA, and in rank order across the cluster through
// distributed shared memory; rank 0 solves the 6x6 system, applies exp(update) * T21, takes the
// reference's termination decisions and publishes the next pose to the other CTAs through its
// shared memory.  No host round trip, no global-memory traffic and no atomics between iterations;
// batched frames are independent clusters of the same launch.  Images are read through the
// read-only path (__ldg): level 0 of a 640x480 frame is 300 kB and stays in L1/L2.
#include "ba_math.cuh"
#include "image_math.cuh"
#include "../../include/viso_direct.h"

#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

namespace cg = cooperative_groups;

namespace visodirect {

using visoimg::DevImage;
using visoimg::is_inside;
using visoimg::pixel_value;

constexpr int kClusterMax = 16;  // CTAs per frame when few frames run: the largest (non-portable) cluster a B200 schedules
constexpr int kClusterBatch = 8; // ... and when many frames run: the portable size packs the GPCs better
constexpr int kThreads = 512;  // threads per CTA
constexpr int kWarps = kThreads / 32;
constexpr int kSums = 28;      // 21 (upper triangle of H) + 6 (b) + 1 (cost)
constexpr int kLevels = VISO_DIRECT_MAX_LEVELS;

struct DevFrame {
    DevImage ref[kLevels], cur[kLevels];
    const double* points;
    int32_t n_points, pad;
    double ref_pose[7];
};

struct DevParams {
    double fx, fy, cx, cy, delta_thresh;
    int32_t hp, max_iterations, reference_termination, pad;
};

// ---- device arithmetic, statement by statement as the reference writes it ----

// keyframe.h:82-89 with R|t already expanded (rt[0..8] = R row-major, rt[9..11] = t)
__device__ __forceinline__ void project(const DevParams& K, const double* rt, double X, double Y, double Z,
    double scale, double& u, double& v, double& x, double& y, double& z)
{
    x = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
    y = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
    z = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
    const double xn = x / z, yn = y / z;
    u = scale * (xn * K.fx + K.cx);
    v = scale * (yn * K.fy + K.cy);
}

// update = H.inverse() * b (viso.cpp:735): Gauss-Jordan with partial pivoting on [H | I]
__device__ void inverse_times(const double* H, const double* b, double* x)
{
    double A[6][12];
    for (int r = 0; r < 6; ++r)
        for (int q = 0; q < 6; ++q) A[r][q] = H[6 * r + q], A[r][6 + q] = r == q ? 1.0 : 0.0;
    for (int k = 0; k < 6; ++k) {
        int p = k;
        for (int r = k + 1; r < 6; ++r)
            if (fabs(A[r][k]) > fabs(A[p][k])) p = r;
        if (p != k)
            for (int q = 0; q < 12; ++q) {
                const double t = A[k][q];
                A[k][q] = A[p][q], A[p][q] = t;
            }
        const double inv = 1.0 / A[k][k];
        for (int q = 0; q < 12; ++q) A[k][q] *= inv;
        for (int r = 0; r < 6; ++r) {
            if (r == k) continue;
            const double f = A[r][k];
            for (int q = 0; q < 12; ++q) A[r][q] -= f * A[k][q];
        }
    }
    for (int r = 0; r < 6; ++r) {
        double s = 0.0;
        for (int q = 0; q < 6; ++q) s += A[r][6 + q] * b[q];
        x[r] = s;
    }
}

// Gauss-Newton state of one frame on one level: lives in rank 0's shared memory, touched by its thread 0
struct GnState {
    double T21[7], best[7];
    double cost, lastCost;
    int applied, reason, nGood, pad;
};

// viso.cpp:735-757: solve, update, and the reference's termination decisions.  Returns 1 when the
// level is finished.  Kept out of line so that its 6x12 elimination tableau does not cost the
// accumulation loop its registers.
__device__ __noinline__ int gn_decide(GnState& st, const double* tot, const DevParams& K, int iter)
{
    double H[36], b[6], update[6];
    int e = 0;
    for (int r = 0; r < 6; ++r)
        for (int q = r; q < 6; ++q) H[6 * r + q] = H[6 * q + r] = tot[e++];
    for (int q = 0; q < 6; ++q) b[q] = tot[21 + q];
    st.nGood = (int)tot[kSums];
    st.cost = (K.reference_termination ? st.cost : 0.0) + tot[27]; // never reset in the reference (viso.cpp:673, :729)
    inverse_times(H, b, update);                  // viso.cpp:735
    double next[7];
    visoba::pose_oplus(update, st.T21, next);     // viso.cpp:737
    for (int k = 0; k < 7; ++k) st.T21[k] = next[k];
    ++st.applied;
    st.cost /= st.nGood;                          // viso.cpp:739
    if (isnan(update[0])) {                       // viso.cpp:741-744
        for (int k = 0; k < 7; ++k) st.T21[k] = st.best[k];
        st.reason = 1;
        return 1;
    }
    if (iter > 0 && st.cost > st.lastCost) {      // viso.cpp:746-749
        for (int k = 0; k < 7; ++k) st.T21[k] = st.best[k];
        st.reason = 2;
        return 1;
    }
    if ((K.reference_termination || iter > 0) && (1 - st.cost / st.lastCost) < K.delta_thresh) {
        st.reason = 3;                            // viso.cpp:751-753
        return 1;
    }
    for (int k = 0; k < 7; ++k) st.best[k] = st.T21[k]; // viso.cpp:755-756
    st.lastCost = st.cost;
    return 0;
}

// ------------------------------------------------------------------------------------------
// One cluster per frame: Gauss-Newton over the levels level_hi .. level_lo (viso.cpp:661-766),
// or (linearize_only) the sums of one linearisation at the given pose.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
k_direct(const DevFrame* __restrict__ frames, const DevParams K, double* __restrict__ poses,
    viso_direct_level_stats* __restrict__ stats, int level_hi, int level_lo, int linearize_only,
    double* __restrict__ lin_out)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank(), n_ranks = (int)cluster.num_blocks();
    const int f = blockIdx.x / n_ranks;
    const DevFrame& F = frames[f];

    __shared__ double s_warp[kWarps][kSums + 1]; // [.][kSums] = good points (exact in a double)
    __shared__ double s_part[kSums + 1];         // this CTA's sums; read by rank 0 through DSMEM
    __shared__ double s_tot[kSums + 1];
    __shared__ GnState s_gn;                     // rank 0: s_gn.T21 is the pose every CTA linearises at
    __shared__ int s_flag;                       // rank 0: 1 = the level is finished
    __shared__ double s_rt[12], s_ref_rt[12];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gw = rank * kWarps + warp, nw = n_ranks * kWarps;
    const int side = 2 * K.hp, npix = side * side;
    const double hp = (double)K.hp;

    if (rank == 0 && tid == 0) {
        for (int k = 0; k < 7; ++k) s_gn.T21[k] = poses[7 * (size_t)f + k];
        s_flag = 0;
    }
    if (tid == 0) visoba::write_cam_rt(F.ref_pose, s_ref_rt);
    cluster.sync();

    for (int level = level_hi; level >= level_lo; --level) {
        const double scale = 1.0 / (double)(1 << level); // 1, 0.5, 0.25, 0.125 (viso.cpp:666)
        const DevImage ref = F.ref[level], cur = F.cur[level];
        if (rank == 0 && tid == 0) {
            s_gn.cost = 0, s_gn.lastCost = 0; // viso.cpp:673
            s_gn.applied = 0, s_gn.reason = 0, s_gn.nGood = 0;
            for (int k = 0; k < 7; ++k) s_gn.best[k] = s_gn.T21[k]; // viso.cpp:676
        }

        for (int iter = 0; iter < K.max_iterations; ++iter) {
            { // the pose published by rank 0 (current_frame->SetR / SetT, viso.cpp:685-686)
                const double* rp = cluster.map_shared_rank(s_gn.T21, 0);
                if (tid == 0) {
                    double p[7];
                    for (int k = 0; k < 7; ++k) p[k] = rp[k];
                    visoba::write_cam_rt(p, s_rt);
                }
                __syncthreads();
            }

            double acc[kSums];
#pragma unroll
            for (int k = 0; k < kSums; ++k) acc[k] = 0.0;
            int good_cnt = 0;
            // The warp's points come in batches of 32: lane l prepares point gw + (batch * 32 + l) * nw (projections,
            // the inside test, dPixeldXi: 14 fp64 divisions that would otherwise be repeated by all 32 lanes), then the
            // good points are taken in order, their 14 values broadcast by shuffles, and the lanes split the patch.
            for (int base = gw; base < F.n_points; base += 32 * nw) { // viso.cpp:688
                const int i = base + lane * nw;
                double q_ur = 0, q_vr = 0, q_uc = 0, q_vc = 0, q_j[10];
#pragma unroll
                for (int q = 0; q < 10; ++q) q_j[q] = 0.0;
                bool good = false;
                if (i < F.n_points) {
                    const double X = __ldg(F.points + 3 * (size_t)i), Y = __ldg(F.points + 3 * (size_t)i + 1);
                    const double Z = __ldg(F.points + 3 * (size_t)i + 2);
                    double x, y, z, xr, yr, zr;
                    project(K, s_ref_rt, X, Y, Z, scale, q_ur, q_vr, xr, yr, zr); // viso.cpp:693-695
                    project(K, s_rt, X, Y, Z, scale, q_uc, q_vc, x, y, z);         // viso.cpp:697-699
                    good = is_inside(ref, q_ur - hp, q_vr - hp) && is_inside(ref, q_ur + hp, q_vr + hp)
                        && is_inside(cur, q_uc - hp, q_vc - hp) && is_inside(cur, q_uc + hp, q_vc + hp); // viso.cpp:704-715
                    // dPixeldXi (viso.cpp:640-658); its two structural zeros are not carried
                    const double fx = K.fx * scale, fy = K.fy * scale, zz = z * z, xy = x * y;
                    q_j[0] = fx / z, q_j[1] = -fx * x / zz, q_j[2] = -fx * xy / zz, q_j[3] = fx + fx * x * x / zz, q_j[4] = -fx * y / z;
                    q_j[5] = fy / z, q_j[6] = -fy * y / zz, q_j[7] = -fy - fy * y * y / zz, q_j[8] = fy * xy / zz, q_j[9] = fy * x / z;
                }
                unsigned todo = __ballot_sync(0xffffffffu, good);
                good_cnt += __popc(todo);
                while (todo) {
                    const int src = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const double u_ref = __shfl_sync(0xffffffffu, q_ur, src), v_ref = __shfl_sync(0xffffffffu, q_vr, src);
                    const double u_cur = __shfl_sync(0xffffffffu, q_uc, src), v_cur = __shfl_sync(0xffffffffu, q_vc, src);
                    double J0[6], J1[6];
                    J0[0] = __shfl_sync(0xffffffffu, q_j[0], src), J0[1] = 0;
                    J1[0] = 0, J1[1] = __shfl_sync(0xffffffffu, q_j[5], src);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        J0[2 + q] = __shfl_sync(0xffffffffu, q_j[1 + q], src);
                        J1[2 + q] = __shfl_sync(0xffffffffu, q_j[6 + q], src);
                    }
                for (int k = lane; k < npix; k += 32) { // viso.cpp:722-723: x outer, y inner
                    const double px = (double)(k / side - K.hp), py = (double)(k % side - K.hp);
                    const double uc = u_cur + px, vc = v_cur + py;
                    const double error = pixel_value(ref, u_ref + px, v_ref + py) - pixel_value(cur, uc, vc);
                    // GetGradient, keyframe.h:59-64
                    const double g0 = 0.5 * (pixel_value(cur, uc + 1, vc) - pixel_value(cur, uc - 1, vc));
                    const double g1 = 0.5 * (pixel_value(cur, uc, vc + 1) - pixel_value(cur, uc, vc - 1));
                    double J[6];
#pragma unroll
                    for (int q = 0; q < 6; ++q) J[q] = -(g0 * J0[q] + g1 * J1[q]); // viso.cpp:726
                    int e = 0;
#pragma unroll
                    for (int r = 0; r < 6; ++r)
#pragma unroll
                        for (int q = r; q < 6; ++q) acc[e++] += J[r] * J[q];   // viso.cpp:727 (upper triangle)
#pragma unroll
                    for (int q = 0; q < 6; ++q) acc[21 + q] += -error * J[q];  // viso.cpp:728
                    acc[27] += error * error;                                  // viso.cpp:729
                }
                }
            }

            // fixed-shape sum: lanes (butterfly) -> warps in order -> CTAs in rank order
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1)
#pragma unroll
                for (int k = 0; k < kSums; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], off);
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < kSums; ++k) s_warp[warp][k] = acc[k];
                s_warp[warp][kSums] = (double)good_cnt;
            }
            __syncthreads();
            if (tid <= kSums) {
                double s = 0.0;
                for (int w = 0; w < kWarps; ++w) s += s_warp[w][tid];
                s_part[tid] = s;
            }
            cluster.sync();
            if (rank == 0) {
                if (tid <= kSums) {
                    double s = 0.0;
                    for (int r = 0; r < n_ranks; ++r) s += cluster.map_shared_rank(s_part, r)[tid];
                    s_tot[tid] = s;
                }
                __syncthreads();
                if (tid == 0) {
                    if (linearize_only) {
                        double* o = lin_out + 44 * (size_t)f;
                        int e = 0;
                        for (int r = 0; r < 6; ++r)
                            for (int q = r; q < 6; ++q) o[6 * r + q] = o[6 * q + r] = s_tot[e++];
                        for (int q = 0; q < 6; ++q) o[36 + q] = s_tot[21 + q];
                        o[42] = s_tot[27], o[43] = s_tot[kSums];
                        s_flag = 1;
                    } else {
                        s_flag = gn_decide(s_gn, s_tot, K, iter);
                    }
                }
            }
            cluster.sync();
            if (*cluster.map_shared_rank(&s_flag, 0)) break;
        }
        if (rank == 0 && tid == 0 && stats && !linearize_only) {
            viso_direct_level_stats& st = stats[kLevels * (size_t)f + level];
            st.iterations = s_gn.applied, st.n_good = s_gn.nGood, st.stop_reason = s_gn.reason, st.reserved = 0;
            st.cost = s_gn.cost, st.last_cost = s_gn.lastCost;
        }
        if (linearize_only) break;
    }
    if (rank == 0 && tid == 0 && !linearize_only)
        for (int k = 0; k < 7; ++k) poses[7 * (size_t)f + k] = s_gn.T21[k];
    cluster.sync(); // nobody leaves while its shared memory may still be read by a peer
}

} // namespace visodirect

// ------------------------------------------------------------------------------------------
// host side of the C ABI
// ------------------------------------------------------------------------------------------
using namespace visodirect;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct viso_direct {
    viso_direct_config cfg;
    std::string err;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool usable = false;
    int64_t launches = 0;
    std::vector<DevFrame> frames;       // host mirror of the device descriptors
    std::vector<DevBuf> images;         // [frame][which][level]
    std::vector<DevBuf> frame_points;   // per-frame point sets
    std::vector<int32_t> frame_npts;    // -1: the frame uses the shared set
    DevBuf shared_points;
    int32_t shared_npts = 0;
    DevBuf d_frames, d_poses, d_stats, d_lin;
    bool frames_dirty = true;
};

namespace {

int32_t fail(viso_direct* h, int32_t code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    return code;
}

#define DCU(call)                                                                                   \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess)                                                                      \
            return fail(h, VISO_DIRECT_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                __FILE__, __LINE__);                                                                \
    } while (0)

int32_t reserve(viso_direct* h, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return VISO_DIRECT_OK;
    if (b.p) DCU(cudaFree(b.p));
    b.p = nullptr, b.cap = 0;
    DCU(cudaMalloc(&b.p, bytes));
    b.cap = bytes;
    h->frames_dirty = true;
    return VISO_DIRECT_OK;
}

int32_t need_device(viso_direct* h)
{
    if (!h) return VISO_DIRECT_ERR_ARG;
    if (!h->usable)
        return fail(h, VISO_DIRECT_ERR_CUDA, "no usable sm_100 device: %s (this backend has no CPU fallback)",
            h->err.empty() ? "not initialised" : h->err.c_str());
    DCU(cudaSetDevice(h->cfg.device));
    return VISO_DIRECT_OK;
}

bool finite_all(const double* p, size_t n)
{
    for (size_t i = 0; i < n; ++i)
        if (!std::isfinite(p[i])) return false;
    return true;
}

int32_t init_device(viso_direct* h)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(h, VISO_DIRECT_ERR_CUDA, "no CUDA device usable (%s)", e == cudaSuccess ? "0 devices" : cudaGetErrorString(e));
    if (h->cfg.device < 0 || h->cfg.device >= n)
        return fail(h, VISO_DIRECT_ERR_CUDA, "device %d out of range (%d devices)", h->cfg.device, n);
    cudaDeviceProp prop;
    DCU(cudaGetDeviceProperties(&prop, h->cfg.device));
    if (prop.major != 10)
        return fail(h, VISO_DIRECT_ERR_CUDA, "device %d is sm_%d%d; the kernels are built for sm_100a only",
            h->cfg.device, prop.major, prop.minor);
    DCU(cudaSetDevice(h->cfg.device));
    DCU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    DCU(cudaEventCreate(&h->ev0));
    DCU(cudaEventCreate(&h->ev1));
    h->usable = true;
    h->err.clear();
    return VISO_DIRECT_OK;
}

// descriptors of every frame -> device (only when something changed)
int32_t sync_frames(viso_direct* h, int32_t n_frames)
{
    for (int32_t f = 0; f < n_frames; ++f) {
        DevFrame& F = h->frames[f];
        const bool own = h->frame_npts[f] >= 0;
        const double* pts = (const double*)(own ? h->frame_points[f].p : h->shared_points.p);
        const int32_t np = own ? h->frame_npts[f] : h->shared_npts;
        if (F.points != pts || F.n_points != np) F.points = pts, F.n_points = np, h->frames_dirty = true;
    }
    if (!h->frames_dirty) return VISO_DIRECT_OK;
    int32_t s = reserve(h, h->d_frames, sizeof(DevFrame) * h->frames.size());
    if (s) return s;
    DCU(cudaMemcpyAsync(h->d_frames.p, h->frames.data(), sizeof(DevFrame) * h->frames.size(), cudaMemcpyHostToDevice, h->stream));
    h->frames_dirty = false;
    return VISO_DIRECT_OK;
}

int32_t check_ready(viso_direct* h, int32_t n_frames, int32_t level_hi, int32_t level_lo)
{
    if (n_frames < 1 || n_frames > h->cfg.max_frames) return fail(h, VISO_DIRECT_ERR_ARG, "n_frames %d outside 1..%d", n_frames, h->cfg.max_frames);
    for (int32_t f = 0; f < n_frames; ++f) {
        for (int l = level_lo; l <= level_hi; ++l)
            if (!h->frames[f].ref[l].data || !h->frames[f].cur[l].data)
                return fail(h, VISO_DIRECT_ERR_STATE, "frame %d level %d: image not set", f, l);
        const int32_t np = h->frame_npts[f] >= 0 ? h->frame_npts[f] : h->shared_npts;
        if (np <= 0) return fail(h, VISO_DIRECT_ERR_STATE, "frame %d has no map points", f);
    }
    return VISO_DIRECT_OK;
}

DevParams params_of(const viso_direct_config& c)
{
    DevParams K;
    K.fx = c.fx, K.fy = c.fy, K.cx = c.cx, K.cy = c.cy, K.delta_thresh = c.delta_thresh;
    K.hp = c.half_patch_size, K.max_iterations = c.max_iterations, K.reference_termination = c.reference_termination, K.pad = 0;
    return K;
}

} // namespace

extern "C" {

void viso_direct_default_config(viso_direct_config* c)
{
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->struct_size = (int32_t)sizeof(*c);
    c->device = 0;
    c->max_points = 4096;
    c->max_frames = 1;
    c->max_width = 640, c->max_height = 480;                    // src/main.cpp:73
    c->fx = 517.3, c->fy = 516.5, c->cx = 325.1, c->cy = 249.7; // src/main.cpp:14-17
    c->half_patch_size = 4;                                     // include/viso.h:25
    c->max_iterations = 100;                                    // src/viso.cpp:671
    c->delta_thresh = 0.005;                                    // src/viso.cpp:668
    c->reference_termination = 1;
}

viso_direct* viso_direct_create(const viso_direct_config* cfg, int32_t* status)
{
    int32_t dummy;
    if (!status) status = &dummy;
    if (!cfg || cfg->struct_size != (int32_t)sizeof(viso_direct_config) || cfg->max_points < 1 || cfg->max_frames < 1
        || cfg->half_patch_size < 1 || cfg->half_patch_size > 16 || cfg->max_iterations < 1) {
        *status = VISO_DIRECT_ERR_ARG;
        return nullptr;
    }
    viso_direct* h = new viso_direct();
    h->cfg = *cfg;
    h->frames.assign(cfg->max_frames, DevFrame());
    for (auto& F : h->frames) {
        memset(&F, 0, sizeof(F));
        F.ref_pose[3] = 1.0; // identity: frame 0 of viso sits at the origin (keyframe.h:34-35)
    }
    h->images.assign((size_t)cfg->max_frames * 2 * kLevels, DevBuf());
    h->frame_points.assign(cfg->max_frames, DevBuf());
    h->frame_npts.assign(cfg->max_frames, -1);
    *status = init_device(h); // a handle without a device still answers last_error(); compute calls fail
    return h;
}

void viso_direct_destroy(viso_direct* h)
{
    if (!h) return;
    if (h->usable) {
        cudaSetDevice(h->cfg.device);
        cudaStreamSynchronize(h->stream);
        for (auto& b : h->images) cudaFree(b.p);
        for (auto& b : h->frame_points) cudaFree(b.p);
        cudaFree(h->shared_points.p), cudaFree(h->d_frames.p), cudaFree(h->d_poses.p), cudaFree(h->d_stats.p), cudaFree(h->d_lin.p);
        cudaEventDestroy(h->ev0), cudaEventDestroy(h->ev1);
        cudaStreamDestroy(h->stream);
    }
    delete h;
}

const char* viso_direct_last_error(const viso_direct* h) { return h ? h->err.c_str() : "null handle"; }
int64_t viso_direct_kernel_launches(const viso_direct* h) { return h ? h->launches : 0; }

int32_t viso_direct_set_image(viso_direct* h, int32_t frame, int32_t which, int32_t level,
                              const uint8_t* data, int32_t cols, int32_t rows, int32_t step)
{
    int32_t s = need_device(h);
    if (s) return s;
    if (!data || frame < 0 || frame >= h->cfg.max_frames || which < 0 || which > 1 || level < 0 || level >= kLevels
        || cols < 1 || rows < 1 || step < cols)
        return fail(h, VISO_DIRECT_ERR_ARG, "set_image: bad argument (frame %d which %d level %d %dx%d step %d)", frame, which, level, cols, rows, step);
    if (cols > h->cfg.max_width || rows > h->cfg.max_height)
        return fail(h, VISO_DIRECT_ERR_CAPACITY, "set_image: %dx%d exceeds the handle's %dx%d", cols, rows, h->cfg.max_width, h->cfg.max_height);
    DevBuf& b = h->images[((size_t)frame * 2 + which) * kLevels + level];
    const size_t bytes = (size_t)rows * step;
    if ((s = reserve(h, b, bytes))) return s;
    DCU(cudaMemcpyAsync(b.p, data, bytes, cudaMemcpyHostToDevice, h->stream));
    DCU(cudaStreamSynchronize(h->stream)); // the caller may reuse `data` on return
    DevImage& im = which == 0 ? h->frames[frame].ref[level] : h->frames[frame].cur[level];
    if (im.data != b.p || im.cols != cols || im.rows != rows || im.step != step) h->frames_dirty = true;
    im.data = (const uint8_t*)b.p, im.cols = cols, im.rows = rows, im.step = step, im.pad = 0;
    return VISO_DIRECT_OK;
}

int32_t viso_direct_set_frame(viso_direct* h, int32_t frame, int32_t which, const uint8_t* data, int32_t cols, int32_t rows, int32_t step)
{
    int32_t s = viso_direct_set_image(h, frame, which, 0, data, cols, rows, step);
    if (s) return s;
    for (int l = 1; l < kLevels; ++l) { // keyframe.h:40-45
        DevImage* pyr = which == 0 ? h->frames[frame].ref : h->frames[frame].cur;
        const DevImage prev = pyr[l - 1];
        const int dcols = (int)(prev.cols * 0.5), drows = (int)(prev.rows * 0.5);
        if (dcols < 1 || drows < 1) return fail(h, VISO_DIRECT_ERR_ARG, "set_frame: %dx%d is too small for %d pyramid levels", cols, rows, kLevels);
        DevBuf& b = h->images[((size_t)frame * 2 + which) * kLevels + l];
        if ((s = reserve(h, b, (size_t)drows * dcols))) return s;
        h->launches += visoimg::launch_pyr_down(prev.data, prev.cols, prev.rows, prev.step, (uint8_t*)b.p, dcols, drows, dcols, h->stream);
        DCU(cudaGetLastError());
        pyr[l].data = (const uint8_t*)b.p, pyr[l].cols = dcols, pyr[l].rows = drows, pyr[l].step = dcols, pyr[l].pad = 0;
    }
    h->frames_dirty = true;
    DCU(cudaStreamSynchronize(h->stream));
    return VISO_DIRECT_OK;
}

int32_t viso_direct_set_points(viso_direct* h, int32_t frame, const double* xyz, int32_t n)
{
    int32_t s = need_device(h);
    if (s) return s;
    if (!xyz || n < 1 || frame >= h->cfg.max_frames) return fail(h, VISO_DIRECT_ERR_ARG, "set_points: bad argument");
    if (n > h->cfg.max_points) return fail(h, VISO_DIRECT_ERR_CAPACITY, "set_points: %d points exceed the pool of %d", n, h->cfg.max_points);
    if (!finite_all(xyz, 3 * (size_t)n)) return fail(h, VISO_DIRECT_ERR_NONFINITE, "set_points: NaN/Inf in the map points");
    DevBuf& b = frame < 0 ? h->shared_points : h->frame_points[frame];
    if ((s = reserve(h, b, sizeof(double) * 3 * (size_t)h->cfg.max_points))) return s;
    DCU(cudaMemcpyAsync(b.p, xyz, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, h->stream));
    DCU(cudaStreamSynchronize(h->stream));
    if (frame < 0) h->shared_npts = n;
    else h->frame_npts[frame] = n;
    return VISO_DIRECT_OK;
}

int32_t viso_direct_set_reference_pose(viso_direct* h, int32_t frame, const double pose[7])
{
    if (!h) return VISO_DIRECT_ERR_ARG;
    if (!pose || frame < 0 || frame >= h->cfg.max_frames) return fail(h, VISO_DIRECT_ERR_ARG, "set_reference_pose: bad argument");
    if (!finite_all(pose, 7)) return fail(h, VISO_DIRECT_ERR_NONFINITE, "set_reference_pose: NaN/Inf");
    memcpy(h->frames[frame].ref_pose, pose, 7 * sizeof(double));
    h->frames_dirty = true;
    return VISO_DIRECT_OK;
}

// cluster launch: 16 CTAs per cluster is beyond the portable size of 8 and has to be allowed explicitly
static cudaError_t launch_direct(viso_direct* h, int32_t n_frames, double* d_poses, viso_direct_level_stats* d_stats,
                                 int level_hi, int level_lo, int linearize_only, double* d_lin)
{
    static visoba::PerDeviceOnce once; // per device: the attribute applies to the current device only
    if (!once.run([] { return cudaFuncSetAttribute(k_direct, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess; }))
        return cudaErrorInvalidValue;
    const int csize = n_frames >= 2 * kClusterBatch ? kClusterBatch : kClusterMax;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(csize * n_frames), lc.blockDim = dim3(kThreads), lc.dynamicSmemBytes = 0, lc.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = csize, at[0].val.clusterDim.y = 1, at[0].val.clusterDim.z = 1;
    lc.attrs = at, lc.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&lc, k_direct, (const DevFrame*)h->d_frames.p, params_of(h->cfg), d_poses, d_stats,
        level_hi, level_lo, linearize_only, d_lin);
    if (e != cudaSuccess) return e;
    ++h->launches;
    return cudaGetLastError();
}

static int32_t run(viso_direct* h, int32_t level_hi, int32_t level_lo, int32_t n_frames, double* poses,
                   viso_direct_level_stats* stats, float* device_ms)
{
    int32_t s = need_device(h);
    if (s) return s;
    if (!poses || level_lo < 0 || level_hi >= kLevels || level_lo > level_hi) return fail(h, VISO_DIRECT_ERR_ARG, "estimate: bad argument");
    if ((s = check_ready(h, n_frames, level_hi, level_lo))) return s;
    if (!finite_all(poses, 7 * (size_t)n_frames)) return fail(h, VISO_DIRECT_ERR_NONFINITE, "estimate: NaN/Inf in the initial poses");
    if ((s = sync_frames(h, n_frames))) return s;
    if ((s = reserve(h, h->d_poses, sizeof(double) * 7 * (size_t)h->cfg.max_frames))) return s;
    if ((s = reserve(h, h->d_stats, sizeof(viso_direct_level_stats) * kLevels * (size_t)h->cfg.max_frames))) return s;
    DCU(cudaMemsetAsync(h->d_stats.p, 0, sizeof(viso_direct_level_stats) * kLevels * (size_t)n_frames, h->stream));
    DCU(cudaMemcpyAsync(h->d_poses.p, poses, sizeof(double) * 7 * (size_t)n_frames, cudaMemcpyHostToDevice, h->stream));
    DCU(cudaEventRecord(h->ev0, h->stream));
    DCU(launch_direct(h, n_frames, (double*)h->d_poses.p, (viso_direct_level_stats*)h->d_stats.p, level_hi, level_lo, 0, nullptr));
    DCU(cudaEventRecord(h->ev1, h->stream));
    DCU(cudaMemcpyAsync(poses, h->d_poses.p, sizeof(double) * 7 * (size_t)n_frames, cudaMemcpyDeviceToHost, h->stream));
    std::vector<viso_direct_level_stats> st;
    if (stats) {
        st.resize((size_t)kLevels * n_frames);
        DCU(cudaMemcpyAsync(st.data(), h->d_stats.p, sizeof(viso_direct_level_stats) * st.size(), cudaMemcpyDeviceToHost, h->stream));
    }
    DCU(cudaStreamSynchronize(h->stream));
    if (device_ms) DCU(cudaEventElapsedTime(device_ms, h->ev0, h->ev1));
    if (stats) {
        if (level_hi == level_lo) // single layer: one entry per frame
            for (int32_t f = 0; f < n_frames; ++f) stats[f] = st[(size_t)kLevels * f + level_hi];
        else
            memcpy(stats, st.data(), sizeof(viso_direct_level_stats) * st.size());
    }
    return VISO_DIRECT_OK;
}

int32_t viso_direct_estimate_level(viso_direct* h, int32_t level, int32_t n_frames, double* poses,
                                   viso_direct_level_stats* stats)
{
    return run(h, level, level, n_frames, poses, stats, nullptr);
}

int32_t viso_direct_estimate(viso_direct* h, int32_t n_frames, double* poses,
                             viso_direct_level_stats* stats, float* device_ms)
{
    return run(h, kLevels - 1, 0, n_frames, poses, stats, device_ms);
}

int32_t viso_direct_linearize(viso_direct* h, int32_t level, int32_t n_frames, const double* poses,
                              double* H, double* b, double* cost, int32_t* n_good)
{
    int32_t s = need_device(h);
    if (s) return s;
    if (!poses || !H || !b || !cost || !n_good || level < 0 || level >= kLevels) return fail(h, VISO_DIRECT_ERR_ARG, "linearize: bad argument");
    if ((s = check_ready(h, n_frames, level, level))) return s;
    if ((s = sync_frames(h, n_frames))) return s;
    if ((s = reserve(h, h->d_poses, sizeof(double) * 7 * (size_t)h->cfg.max_frames))) return s;
    if ((s = reserve(h, h->d_lin, sizeof(double) * 44 * (size_t)h->cfg.max_frames))) return s;
    DCU(cudaMemcpyAsync(h->d_poses.p, poses, sizeof(double) * 7 * (size_t)n_frames, cudaMemcpyHostToDevice, h->stream));
    DCU(launch_direct(h, n_frames, (double*)h->d_poses.p, nullptr, level, level, 1, (double*)h->d_lin.p));
    std::vector<double> out(44 * (size_t)n_frames);
    DCU(cudaMemcpyAsync(out.data(), h->d_lin.p, sizeof(double) * out.size(), cudaMemcpyDeviceToHost, h->stream));
    DCU(cudaStreamSynchronize(h->stream));
    for (int32_t f = 0; f < n_frames; ++f) {
        memcpy(H + 36 * (size_t)f, out.data() + 44 * (size_t)f, 36 * sizeof(double));
        memcpy(b + 6 * (size_t)f, out.data() + 44 * (size_t)f + 36, 6 * sizeof(double));
        cost[f] = out[44 * (size_t)f + 42];
        n_good[f] = (int32_t)out[44 * (size_t)f + 43];
    }
    return VISO_DIRECT_OK;
}

} // extern "C"
