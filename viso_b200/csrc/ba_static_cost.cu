// ba_static_cost.cu — viso_ba_eval_cost_static (include/viso_ba.h): the robustified reprojection cost of a raw edge list
// with NO allocation of any kind.  No counterpart in the reference: added for callers
// that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It is the eval-cost pass of the general path (k_cam_rt + k_eval_cost + the chunk / problem sums, ba_kernels.cu; the same
// obs_geometry and Huber kernel from ba_math.cuh) with the two pieces of scratch that pass needs — the R|t table of every
// pose and the partial sums of the CTAs — held in ONE file-scope `static __device__` array instead of the handle's arena:
//
//     k_static_cam_rt    pose (q xyzw, t)  ->  scratch.cam_rt[pose][12]
//     k_static_cost      CTA c sums the edges [c * per, (c + 1) * per)  ->  scratch.partial[c]
//     k_static_fold      one CTA adds the partials in index order       ->  *cost_out
//
// Every summation has a fixed shape (thread: its edges in order; warp: butterfly; CTA: warps in order; fold: a fixed
// tree over the partials in index order), so two calls on the same input give the same bits.  The order differs from the
// handle's pass (which sums per point first), so the two entry points agree to rounding (~1e-15 relative), not to the bit.
//
// The scratch array:
//   - sizeof(StaticScratch) = 4096 poses x 12 doubles + 1184 partials = 402,688 bytes, held by every GPU that loads the
//     library until the process exits;
//   - in HOST code its name is a host-side stand-in, not a device address: its address is looked up once per device with
//     cudaGetSymbolAddress (that device current) and handed to the kernels as an ordinary pointer;
//   - shared by every call and never cleared: nothing in it has to start at zero, because a call writes each word
//     (cam_rt[0, n_poses), partial[0, n_cta)) before it reads it.  Calls that overlap on one GPU (two streams, two threads)
//     race on it: serialise them.
#include "../../include/viso_ba.h"
#include "ba_math.cuh"

#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cstdarg>
#include <cstdio>
#include <mutex>

namespace {

using namespace visoba;

constexpr int kStaticMaxPoses = VISO_BA_STATIC_MAX_POSES;
constexpr int kStaticCtas = 148 * 8; // one wave of 8 resident CTAs per SM
constexpr int kStaticThreads = 256;
constexpr int kStaticMaxDevices = 64;

struct StaticScratch {
    double cam_rt[(size_t)kStaticMaxPoses * 12];
    double partial[kStaticCtas];
};

static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

__global__ void __launch_bounds__(kStaticThreads) k_static_cam_rt(StaticScratch* __restrict__ S, const double* __restrict__ poses7, int n_poses)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_poses) return;
    double p[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) p[k] = poses7[7 * (size_t)c + k];
    write_cam_rt(p, S->cam_rt + 12 * (size_t)c);
}

__global__ void __launch_bounds__(kStaticThreads) k_static_cost(StaticScratch* __restrict__ S, const double* __restrict__ points3,
    const int32_t* __restrict__ edge_pose, const int32_t* __restrict__ edge_point, const float* __restrict__ edge_uv,
    long long n_edges, long long per_cta, int n_poses, int n_points, double fx, double fy, double cx, double cy, double huber_delta)
{
    __shared__ double s_warp[kStaticThreads / 32];
    DevProblem P; // only the camera model and the robust kernel are read by obs_geometry
    P.fx = fx, P.fy = fy, P.cx = cx, P.cy = cy, P.huber_delta = huber_delta;
    const long long e0 = (long long)blockIdx.x * per_cta, e1 = min(e0 + per_cta, n_edges);
    double acc = 0.0;
    for (long long e = e0 + threadIdx.x; e < e1; e += kStaticThreads) {
        const int cam = __ldg(edge_pose + e), pt = __ldg(edge_point + e);
        if ((unsigned)cam >= (unsigned)n_poses || (unsigned)pt >= (unsigned)n_points) { // the caller's arrays are on the device:
            acc = nan("");                                                              // an index out of range poisons the cost
            continue;
        }
        const double X = __ldg(points3 + 3 * (size_t)pt), Y = __ldg(points3 + 3 * (size_t)pt + 1), Z = __ldg(points3 + 3 * (size_t)pt + 2);
        ObsGeom g;
        obs_geometry(P, S->cam_rt + 12 * (size_t)cam, X, Y, Z, make_float2(__ldg(edge_uv + 2 * e), __ldg(edge_uv + 2 * e + 1)), g);
        acc += g.rho;
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kStaticThreads / 32; ++w) s += s_warp[w];
        S->partial[blockIdx.x] = s;
    }
}

__global__ void __launch_bounds__(kStaticThreads) k_static_fold(const StaticScratch* __restrict__ S, int n_cta, double* __restrict__ cost_out)
{
    __shared__ double s_warp[kStaticThreads / 32];
    double acc = 0.0;
    for (int c = threadIdx.x; c < n_cta; c += kStaticThreads) acc += S->partial[c];
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kStaticThreads / 32; ++w) s += s_warp[w];
        *cost_out = s;
    }
}

int32_t report(char* err, int32_t err_cap, int32_t code, const char* fmt, ...)
{
    if (err && err_cap > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, (size_t)err_cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

// device address of the scratch array on `device` (current on this thread): looked up once per device
std::mutex g_static_mutex;
StaticScratch* g_static_addr[kStaticMaxDevices] = {};

cudaError_t scratch_address(int device, StaticScratch** out)
{
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_addr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_addr[device] = static_cast<StaticScratch*>(p);
    }
    *out = g_static_addr[device];
    return cudaSuccess;
}

} // namespace

extern "C" int32_t viso_ba_eval_cost_static(void* cuda_stream, int64_t n_poses, const double* poses7_dev, int64_t n_points,
    const double* points3_dev, int64_t n_edges, const int32_t* edge_pose_dev, const int32_t* edge_point_dev, const float* edge_uv_dev,
    double fx, double fy, double cx, double cy, double huber_delta, double* cost_out_dev, char* err, int32_t err_cap)
{
    if (err && err_cap > 0) err[0] = 0;
    if (n_poses < 0 || n_points < 0 || n_edges < 0) return report(err, err_cap, VISO_BA_ERR_ARG, "negative size");
    if (!cost_out_dev) return report(err, err_cap, VISO_BA_ERR_ARG, "cost_out_dev is NULL");
    if ((n_poses > 0 && !poses7_dev) || (n_points > 0 && !points3_dev) || (n_edges > 0 && (!edge_pose_dev || !edge_point_dev || !edge_uv_dev)))
        return report(err, err_cap, VISO_BA_ERR_ARG, "NULL input array");
    if (n_poses > kStaticMaxPoses)
        return report(err, err_cap, VISO_BA_ERR_CAPACITY,
            "viso_ba_eval_cost_static: %lld poses, the static scratch array holds %d (VISO_BA_STATIC_MAX_POSES); use a handle and viso_ba_eval_cost",
            (long long)n_poses, kStaticMaxPoses);
    if (n_points > INT32_MAX || n_edges > VISO_BA_STATIC_MAX_EDGES)
        return report(err, err_cap, VISO_BA_ERR_CAPACITY, "viso_ba_eval_cost_static: %lld points / %lld edges, limits %d / %lld",
            (long long)n_points, (long long)n_edges, INT32_MAX, (long long)VISO_BA_STATIC_MAX_EDGES);
    int device = -1;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return report(err, err_cap, VISO_BA_ERR_CUDA, "cudaGetDevice: %s (this backend has no CPU fallback)", cudaGetErrorString(e));
    if (device < 0 || device >= kStaticMaxDevices) return report(err, err_cap, VISO_BA_ERR_UNSUPPORTED, "device %d outside 0..%d", device, kStaticMaxDevices - 1);
    StaticScratch* S = nullptr;
    if ((e = scratch_address(device, &S)) != cudaSuccess)
        return report(err, err_cap, VISO_BA_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", device, cudaGetErrorString(e));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int n_cta = (int)std::min<int64_t>(kStaticCtas, (n_edges + kStaticThreads - 1) / kStaticThreads);
    if (n_poses > 0) k_static_cam_rt<<<(unsigned)((n_poses + kStaticThreads - 1) / kStaticThreads), kStaticThreads, 0, st>>>(S, poses7_dev, (int)n_poses);
    if (n_cta > 0) {
        const long long per = (n_edges + n_cta - 1) / n_cta;
        k_static_cost<<<n_cta, kStaticThreads, 0, st>>>(S, points3_dev, edge_pose_dev, edge_point_dev, edge_uv_dev,
            (long long)n_edges, per, (int)n_poses, (int)n_points, fx, fy, cx, cy, huber_delta);
    }
    k_static_fold<<<1, kStaticThreads, 0, st>>>(S, n_cta, cost_out_dev);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return report(err, err_cap, VISO_BA_ERR_CUDA, "viso_ba_eval_cost_static: kernel launch on device %d: %s", device, cudaGetErrorString(e));
    return VISO_BA_OK;
}
