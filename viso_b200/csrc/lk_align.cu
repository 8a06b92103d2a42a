// lk_align.cu — B200-native Lucas-Kanade patch alignment of viso's front end (SURVEY.md section 8
// row f4).  C ABI: include/viso_lk.h.
//
// Reference semantics (citations relative to /root/reference):
//   Viso::LKAlignmentSingle       src/viso.cpp:845-925   inverse compositional, 100 iterations, double positions
//   level loop of LKAlignment     src/viso.cpp:828-830
//   Viso::OpticalFlowSingleLevel  src/viso.cpp:259-351   forward or inverse, 10 iterations, cv::Point2f positions
//   Viso::OpticalFlowMultiLevel   src/viso.cpp:353-391
//
// B200 design: every keypoint is an independent 2x2 Gauss-Newton problem on an 8x8 patch, so ONE
// WARP owns a keypoint through all pyramid levels and all iterations: 32 lanes x 2 patch pixels,
// the six sums (H00, H01, H11, b0, b1, cost) folded with a fixed xor butterfly so that every lane
// holds the same totals and takes the reference's branches uniformly — no shared memory, no atomics,
// no host round trip between levels.  In the inverse-compositional case the reference patch, its
// gradient and H do not change between iterations (the reference recomputes the same values): they
// are sampled once per level and kept in registers.  Pyramids stay resident in HBM/L2 and are read
// through the read-only path.
#include "ba_math.cuh"
#include "image_math.cuh"
#include "../../include/viso_lk.h"

#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

namespace visolk {

using visoimg::DevImage;
using visoimg::gradient;
using visoimg::is_inside;
using visoimg::pixel_value;

constexpr int kLevels = VISO_LK_LEVELS;
constexpr int kThreads = 128;    // 4 keypoints per CTA
constexpr int kMaxPerLane = 8;   // patch pixels per lane: half_patch_size <= 8

struct LkParams {
    int32_t hp, align_iterations, flow_iterations, pad;
    double thresh;
};

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// update = H.inverse() * b, Eigen's closed form for 2x2 (1 / det times the adjugate)
__device__ __forceinline__ void solve2(double h00, double h01, double h10, double h11, double b0, double b1, double& u0, double& u1)
{
    const double invdet = 1.0 / (h00 * h11 - h01 * h10);
    const double i00 = h11 * invdet, i10 = -h10 * invdet, i01 = -h01 * invdet, i11 = h00 * invdet;
    u0 = i00 * b0 + i01 * b1;
    u1 = i10 * b0 + i11 * b1;
}

// ------------------------------------------------------------------------------------------
// LKAlignmentSingle over the levels level_hi .. level_lo, one warp per AlignmentPair
// ------------------------------------------------------------------------------------------
template <int kPL> // patch pixels per lane: 2 for the reference's 8x8 patch, kMaxPerLane for larger ones
__global__ void __launch_bounds__(kThreads) k_lk_align(const DevImage* __restrict__ slots, const int32_t* __restrict__ ref_slot,
    int cur_slot, int n, const double* __restrict__ uv_ref, double* __restrict__ uv_cur, uint8_t* __restrict__ success,
    int level_hi, int level_lo, const LkParams P)
{
    const int i = (blockIdx.x * kThreads + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n) return; // whole warps leave together
    const int side = 2 * P.hp, npix = side * side;
    const double hp = (double)P.hp;
    const double urx = uv_ref[2 * (size_t)i], ury = uv_ref[2 * (size_t)i + 1];
    double ucx = uv_cur[2 * (size_t)i], ucy = uv_cur[2 * (size_t)i + 1];
    const int rs = ref_slot[i];
    int succ = 1;
    for (int level = level_hi; level >= level_lo; --level) {
        const double s = 1.0 / (double)(1 << level); // scales[level], viso.cpp:850
        const DevImage ref = slots[kLevels * rs + level], cur = slots[kLevels * cur_slot + level];
        // inverse = true (viso.cpp:848): J = -grad(ref) and the reference patch do not depend on (dx, dy)
        double J0[kPL], J1[kPL], rv[kPL];
        double h00 = 0, h01 = 0, h11 = 0;
#pragma unroll
        for (int q = 0; q < kPL; ++q) {
            const int k = lane + 32 * q;
            J0[q] = J1[q] = rv[q] = 0.0;
            if (k < npix) {
                const int x = k / side - P.hp, y = k % side - P.hp; // viso.cpp:877-878: x outer, y inner
                double gx, gy;
                gradient(ref, urx * s + x, ury * s + y, gx, gy);    // viso.cpp:882-883
                J0[q] = -gx, J1[q] = -gy;
                rv[q] = pixel_value(ref, urx * s + x, ury * s + y);
                h00 += J0[q] * J0[q], h01 += J0[q] * J1[q], h11 += J1[q] * J1[q]; // viso.cpp:888
            }
        }
        h00 = warp_sum(h00), h01 = warp_sum(h01), h11 = warp_sum(h11);
        double dx = 0, dy = 0, cost = 0, lastCost = 0; // viso.cpp:858-860
        succ = 1;
        for (int iter = 0; iter < P.align_iterations; ++iter) {
            if (!is_inside(ref, urx * s + dx - hp, ury * s + dy - hp) || !is_inside(ref, urx * s + dx + hp, ury * s + dy + hp)) {
                succ = 0; // viso.cpp:869-873
                break;
            }
            double b0 = 0, b1 = 0;
            cost = 0;
#pragma unroll
            for (int q = 0; q < kPL; ++q) {
                const int k = lane + 32 * q;
                if (k < npix) {
                    const int x = k / side - P.hp, y = k % side - P.hp;
                    const double error = rv[q] - pixel_value(cur, ucx * s + x + dx, ucy * s + y + dy); // viso.cpp:885
                    b0 += -J0[q] * error, b1 += -J1[q] * error;
                    cost += error * error;
                }
            }
            b0 = warp_sum(b0), b1 = warp_sum(b1), cost = warp_sum(cost);
            double u0, u1;
            solve2(h00, h01, h01, h11, b0, b1, u0, u1); // viso.cpp:894
            if (isnan(u0)) {                            // viso.cpp:896-899
                succ = 0;
                break;
            }
            if (iter > 0 && cost > lastCost) break;     // viso.cpp:900-902
            dx += u0, dy += u1;
            lastCost = cost;
            succ = !(lastCost > P.thresh);              // viso.cpp:909-913
        }
        ucx += dx / s, ucy += dy / s; // viso.cpp:916
    }
    if (lane == 0) {
        uv_cur[2 * (size_t)i] = ucx, uv_cur[2 * (size_t)i + 1] = ucy;
        success[i] = (uint8_t)succ;
    }
}

// ------------------------------------------------------------------------------------------
// OpticalFlowSingleLevel over the levels level_hi .. level_lo (multi: with OpticalFlowMultiLevel's
// rescaling of kp1 / kp2 between levels), one warp per keypoint.  Positions are cv::Point2f.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_lk_flow(const DevImage* __restrict__ slots, int slot1, int slot2, int n,
    const float2* __restrict__ kp1, float2* __restrict__ kp2, int have_initial, int inverse, uint8_t* __restrict__ success,
    int level_hi, int level_lo, int multi, const LkParams P)
{
    const int i = (blockIdx.x * kThreads + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n) return;
    const int side = 2 * P.hp, npix = side * side;
    const double hp = (double)P.hp;
    const float2 k1_0 = kp1[i];
    float2 k2 = have_initial ? kp2[i] : make_float2(0.f, 0.f);
    if (multi && have_initial) // viso.cpp:367-370: Point2f *= double
        k2.x = (float)((double)k2.x * 0.125), k2.y = (float)((double)k2.y * 0.125);
    int succ = 1;
    for (int level = level_hi; level >= level_lo; --level) {
        const double s = 1.0 / (double)(1 << level);
        const DevImage img1 = slots[kLevels * slot1 + level], img2 = slots[kLevels * slot2 + level];
        float2 k1 = k1_0;
        if (multi) k1.x = (float)((double)k1_0.x * s), k1.y = (float)((double)k1_0.y * s); // viso.cpp:373-378
        double dx = 0, dy = 0;
        if (have_initial || level != level_hi) { // viso.cpp:267, :272-275: float differences
            dx = (double)(k2.x - k1.x);
            dy = (double)(k2.y - k1.y);
        }
        double cost = 0, lastCost = 0;
        succ = 1;
        for (int iter = 0; iter < P.flow_iterations; ++iter) {
            if (k1.x + dx <= hp || k1.x + dx >= img1.cols - hp || k1.y + dy <= hp || k1.y + dy >= img1.rows - hp) {
                succ = 0; // viso.cpp:286-289
                break;
            }
            double h00 = 0, h01 = 0, h11 = 0, b0 = 0, b1 = 0;
            cost = 0;
            for (int k = lane; k < npix; k += 32) { // viso.cpp:292-293
                const int x = k / side - P.hp, y = k % side - P.hp;
                const float fx = k1.x + (float)x, fy = k1.y + (float)y; // kp.pt.x + x: a float sum
                double gx, gy;
                if (!inverse) gradient(img2, fx + dx, fy + dy, gx, gy); // viso.cpp:299
                else gradient(img1, (double)fx, (double)fy, gx, gy);    // viso.cpp:302
                const double J0 = -gx, J1 = -gy;
                const double error = pixel_value(img1, (double)fx, (double)fy) - pixel_value(img2, fx + dx, fy + dy); // viso.cpp:305
                h00 += J0 * J0, h01 += J0 * J1, h11 += J1 * J1; // viso.cpp:308-310
                b0 += -J0 * error, b1 += -J1 * error;
                cost += error * error;
            }
            h00 = warp_sum(h00), h01 = warp_sum(h01), h11 = warp_sum(h11);
            b0 = warp_sum(b0), b1 = warp_sum(b1), cost = warp_sum(cost);
            double u0, u1;
            solve2(h00, h01, h01, h11, b0, b1, u0, u1); // viso.cpp:313
            if (isnan(u0)) {                            // viso.cpp:315-321
                succ = 0;
                break;
            }
            if (iter > 0 && cost > lastCost) break;     // viso.cpp:323-325
            dx += u0, dy += u1;
            lastCost = cost;
            succ = !(lastCost > P.thresh);              // viso.cpp:332-336
        }
        k2.x = k1.x + (float)dx, k2.y = k1.y + (float)dy; // viso.cpp:342-349
        if (multi && level != 0)                            // viso.cpp:384-389: Point2f /= pyramid_scale
            k2.x = (float)((double)k2.x / 0.5), k2.y = (float)((double)k2.y / 0.5);
    }
    if (lane == 0) {
        kp2[i] = k2;
        success[i] = (uint8_t)succ;
    }
}

// ------------------------------------------------------------------------------------------
// EdgeDirectProjection::computeError (include/bundle_adjuster.h:70-101): 16 lanes per edge, lane =
// sample idx (j outer, i inner).  The reference stops at the first sample that leaves an image
// (setLevel(1), :91-95): the lanes vote, samples before the first failure are kept, the rest are 0.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_edge_direct_projection(const DevImage* __restrict__ slots, int n,
    const double* __restrict__ points, const int32_t* __restrict__ src_slot, const int32_t* __restrict__ tgt_slot,
    const double* __restrict__ frame_poses, const double* __restrict__ vertex_poses, double fx, double fy, double cx, double cy,
    double* __restrict__ errors, uint8_t* __restrict__ level)
{
    const int t = blockIdx.x * kThreads + threadIdx.x;
    const int e = t >> 4, idx = t & 15;
    const bool live = e < n;
    const int ee = live ? e : n - 1;
    const int ss = src_slot[ee], ts = tgt_slot[ee];
    const DevImage src = slots[kLevels * ss], tgt = slots[kLevels * ts];
    const double X = points[3 * (size_t)ee], Y = points[3 * (size_t)ee + 1], Z = points[3 * (size_t)ee + 2];
    double rt[12];
    visoba::write_cam_rt(vertex_poses + 7 * (size_t)ts, rt); // T->estimate(), :73-75
    const double x = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
    const double y = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
    const double z = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
    const double tu = (fx * x + 0.0 * y + cx * z) / z, tv = (0.0 * x + fy * y + cy * z) / z; // K * (R p + t), / z  (:77-78)
    visoba::write_cam_rt(frame_poses + 7 * (size_t)ss, rt);  // srcFrame->Project(p), :79 (keyframe.h:82-89)
    const double xs = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
    const double ys = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
    const double zs = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
    const double su = 1.0 * ((xs / zs) * fx + cx), sv = 1.0 * ((ys / zs) * fy + cy);
    const int i = (idx & 3) - 2, j = (idx >> 2) - 2; // :83-84
    const float u1 = (float)(tu + (double)i), v1 = (float)(tv + (double)j); // :86-87
    const float u2 = (float)(su + (double)i), v2 = (float)(sv + (double)j); // :88-89
    const bool outside = !is_inside(tgt, u1, v1) || !is_inside(src, u2, v2);
    const unsigned half = (threadIdx.x & 16) ? 0xffff0000u : 0x0000ffffu;
    const unsigned fails = __ballot_sync(0xffffffffu, outside) & half;
    const int first = fails ? (__ffs(fails) - 1) & 15 : 16; // the sample at which the reference aborts
    double err = 0.0;
    if (idx < first) err = pixel_value(src, u2, v2) - pixel_value(tgt, u1, v1); // :97
    if (live) {
        errors[16 * (size_t)e + idx] = err;
        if (idx == 0) level[e] = first < 16 ? 1 : 0;
    }
}

} // namespace visolk

// ------------------------------------------------------------------------------------------
// host side of the C ABI
// ------------------------------------------------------------------------------------------
using namespace visolk;

struct LkBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct viso_lk {
    viso_lk_config cfg;
    std::string err;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool usable = false;
    int64_t launches = 0;
    std::vector<DevImage> slots; // host mirror [max_slots][4]
    std::vector<LkBuf> images;
    bool slots_dirty = true;
    LkBuf d_slots, d_ref_slot, d_uv_ref, d_uv_cur, d_kp1, d_kp2, d_success;
    LkBuf d_pts, d_src, d_tgt, d_fpose, d_vpose, d_err;
};

namespace {

int32_t lk_fail(viso_lk* h, int32_t code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    return code;
}

#define LCU(call)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return lk_fail(h, VISO_LK_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                __FILE__, __LINE__);                                                               \
    } while (0)

int32_t lk_reserve(viso_lk* h, LkBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return VISO_LK_OK;
    if (b.p) LCU(cudaFree(b.p));
    b.p = nullptr, b.cap = 0;
    LCU(cudaMalloc(&b.p, bytes));
    b.cap = bytes;
    return VISO_LK_OK;
}

int32_t lk_need_device(viso_lk* h)
{
    if (!h) return VISO_LK_ERR_ARG;
    if (!h->usable)
        return lk_fail(h, VISO_LK_ERR_CUDA, "no usable sm_100 device: %s (this backend has no CPU fallback)",
            h->err.empty() ? "not initialised" : h->err.c_str());
    LCU(cudaSetDevice(h->cfg.device));
    return VISO_LK_OK;
}

int32_t lk_init_device(viso_lk* h)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return lk_fail(h, VISO_LK_ERR_CUDA, "no CUDA device usable (%s)", e == cudaSuccess ? "0 devices" : cudaGetErrorString(e));
    if (h->cfg.device < 0 || h->cfg.device >= n)
        return lk_fail(h, VISO_LK_ERR_CUDA, "device %d out of range (%d devices)", h->cfg.device, n);
    cudaDeviceProp prop;
    LCU(cudaGetDeviceProperties(&prop, h->cfg.device));
    if (prop.major != 10)
        return lk_fail(h, VISO_LK_ERR_CUDA, "device %d is sm_%d%d; the kernels are built for sm_100a only", h->cfg.device, prop.major, prop.minor);
    LCU(cudaSetDevice(h->cfg.device));
    LCU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    LCU(cudaEventCreate(&h->ev0));
    LCU(cudaEventCreate(&h->ev1));
    h->usable = true;
    h->err.clear();
    return VISO_LK_OK;
}

LkParams lk_params(const viso_lk_config& c)
{
    LkParams P;
    P.hp = c.half_patch_size, P.align_iterations = c.align_iterations, P.flow_iterations = c.flow_iterations, P.pad = 0;
    P.thresh = c.photometric_error_thresh;
    return P;
}

int32_t lk_sync_slots(viso_lk* h)
{
    if (!h->slots_dirty) return VISO_LK_OK;
    int32_t s = lk_reserve(h, h->d_slots, sizeof(DevImage) * h->slots.size());
    if (s) return s;
    LCU(cudaMemcpyAsync(h->d_slots.p, h->slots.data(), sizeof(DevImage) * h->slots.size(), cudaMemcpyHostToDevice, h->stream));
    h->slots_dirty = false;
    return VISO_LK_OK;
}

int32_t lk_check_slot(viso_lk* h, int32_t slot, int32_t level_hi, int32_t level_lo)
{
    if (slot < 0 || slot >= h->cfg.max_slots) return lk_fail(h, VISO_LK_ERR_INDEX, "image slot %d outside 0..%d", slot, h->cfg.max_slots - 1);
    for (int l = level_lo; l <= level_hi; ++l)
        if (!h->slots[(size_t)kLevels * slot + l].data) return lk_fail(h, VISO_LK_ERR_STATE, "slot %d level %d: image not set", slot, l);
    return VISO_LK_OK;
}

bool lk_finite(const double* p, size_t n)
{
    for (size_t i = 0; i < n; ++i)
        if (!std::isfinite(p[i])) return false;
    return true;
}

bool lk_finite_f(const float* p, size_t n)
{
    for (size_t i = 0; i < n; ++i)
        if (!std::isfinite(p[i])) return false;
    return true;
}

int32_t lk_align_run(viso_lk* h, int32_t level_hi, int32_t level_lo, int32_t n, const int32_t* ref_slot, int32_t cur_slot,
                     const double* uv_ref, double* uv_cur, uint8_t* success, float* device_ms)
{
    int32_t s = lk_need_device(h);
    if (s) return s;
    if (n < 0 || (n > 0 && (!ref_slot || !uv_ref || !uv_cur || !success))) return lk_fail(h, VISO_LK_ERR_ARG, "align: bad argument");
    if (n > h->cfg.max_keypoints) return lk_fail(h, VISO_LK_ERR_CAPACITY, "align: %d pairs exceed the pool of %d", n, h->cfg.max_keypoints);
    if (device_ms) *device_ms = 0.f;
    if (n == 0) return VISO_LK_OK; // LKAlignment with no visible map point: nothing to do
    if ((s = lk_check_slot(h, cur_slot, level_hi, level_lo))) return s;
    for (int32_t i = 0; i < n; ++i)
        if ((s = lk_check_slot(h, ref_slot[i], level_hi, level_lo))) return s;
    if (!lk_finite(uv_ref, 2 * (size_t)n) || !lk_finite(uv_cur, 2 * (size_t)n)) return lk_fail(h, VISO_LK_ERR_NONFINITE, "align: NaN/Inf in the pixel positions");
    if ((s = lk_sync_slots(h))) return s;
    const size_t cap = (size_t)h->cfg.max_keypoints;
    if ((s = lk_reserve(h, h->d_ref_slot, sizeof(int32_t) * cap)) || (s = lk_reserve(h, h->d_uv_ref, sizeof(double) * 2 * cap))
        || (s = lk_reserve(h, h->d_uv_cur, sizeof(double) * 2 * cap)) || (s = lk_reserve(h, h->d_success, cap)))
        return s;
    LCU(cudaMemcpyAsync(h->d_ref_slot.p, ref_slot, sizeof(int32_t) * n, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaMemcpyAsync(h->d_uv_ref.p, uv_ref, sizeof(double) * 2 * n, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaMemcpyAsync(h->d_uv_cur.p, uv_cur, sizeof(double) * 2 * n, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaEventRecord(h->ev0, h->stream));
    const int grid = (n * 32 + kThreads - 1) / kThreads;
    if (4 * h->cfg.half_patch_size * h->cfg.half_patch_size <= 64)
        k_lk_align<2><<<grid, kThreads, 0, h->stream>>>((const DevImage*)h->d_slots.p, (const int32_t*)h->d_ref_slot.p, cur_slot, n,
            (const double*)h->d_uv_ref.p, (double*)h->d_uv_cur.p, (uint8_t*)h->d_success.p, level_hi, level_lo, lk_params(h->cfg));
    else
        k_lk_align<kMaxPerLane><<<grid, kThreads, 0, h->stream>>>((const DevImage*)h->d_slots.p, (const int32_t*)h->d_ref_slot.p, cur_slot, n,
            (const double*)h->d_uv_ref.p, (double*)h->d_uv_cur.p, (uint8_t*)h->d_success.p, level_hi, level_lo, lk_params(h->cfg));
    LCU(cudaGetLastError());
    ++h->launches;
    LCU(cudaEventRecord(h->ev1, h->stream));
    LCU(cudaMemcpyAsync(uv_cur, h->d_uv_cur.p, sizeof(double) * 2 * n, cudaMemcpyDeviceToHost, h->stream));
    LCU(cudaMemcpyAsync(success, h->d_success.p, n, cudaMemcpyDeviceToHost, h->stream));
    LCU(cudaStreamSynchronize(h->stream));
    if (device_ms) LCU(cudaEventElapsedTime(device_ms, h->ev0, h->ev1));
    return VISO_LK_OK;
}

int32_t lk_flow_run(viso_lk* h, int32_t level_hi, int32_t level_lo, int32_t multi, int32_t slot1, int32_t slot2, int32_t n,
                    const float* kp1, float* kp2, int32_t have_initial, int32_t inverse, uint8_t* success, float* device_ms)
{
    int32_t s = lk_need_device(h);
    if (s) return s;
    if (n < 0 || (n > 0 && (!kp1 || !kp2 || !success))) return lk_fail(h, VISO_LK_ERR_ARG, "optical_flow: bad argument");
    if (n > h->cfg.max_keypoints) return lk_fail(h, VISO_LK_ERR_CAPACITY, "optical_flow: %d keypoints exceed the pool of %d", n, h->cfg.max_keypoints);
    if (device_ms) *device_ms = 0.f;
    if (n == 0) return VISO_LK_OK;
    if ((s = lk_check_slot(h, slot1, level_hi, level_lo)) || (s = lk_check_slot(h, slot2, level_hi, level_lo))) return s;
    if (!lk_finite_f(kp1, 2 * (size_t)n) || (have_initial && !lk_finite_f(kp2, 2 * (size_t)n)))
        return lk_fail(h, VISO_LK_ERR_NONFINITE, "optical_flow: NaN/Inf in the keypoints");
    if ((s = lk_sync_slots(h))) return s;
    const size_t cap = (size_t)h->cfg.max_keypoints;
    if ((s = lk_reserve(h, h->d_kp1, sizeof(float) * 2 * cap)) || (s = lk_reserve(h, h->d_kp2, sizeof(float) * 2 * cap))
        || (s = lk_reserve(h, h->d_success, cap)))
        return s;
    LCU(cudaMemcpyAsync(h->d_kp1.p, kp1, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, h->stream));
    if (have_initial) LCU(cudaMemcpyAsync(h->d_kp2.p, kp2, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaEventRecord(h->ev0, h->stream));
    const int grid = (n * 32 + kThreads - 1) / kThreads;
    k_lk_flow<<<grid, kThreads, 0, h->stream>>>((const DevImage*)h->d_slots.p, slot1, slot2, n, (const float2*)h->d_kp1.p,
        (float2*)h->d_kp2.p, have_initial, inverse, (uint8_t*)h->d_success.p, level_hi, level_lo, multi, lk_params(h->cfg));
    LCU(cudaGetLastError());
    ++h->launches;
    LCU(cudaEventRecord(h->ev1, h->stream));
    LCU(cudaMemcpyAsync(kp2, h->d_kp2.p, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, h->stream));
    LCU(cudaMemcpyAsync(success, h->d_success.p, n, cudaMemcpyDeviceToHost, h->stream));
    LCU(cudaStreamSynchronize(h->stream));
    if (device_ms) LCU(cudaEventElapsedTime(device_ms, h->ev0, h->ev1));
    return VISO_LK_OK;
}

} // namespace

extern "C" {

void viso_lk_default_config(viso_lk_config* c)
{
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->struct_size = (int32_t)sizeof(*c);
    c->device = 0;
    c->max_keypoints = 4096;
    c->max_slots = 8;
    c->max_width = 640, c->max_height = 480;                  // src/main.cpp:73
    c->half_patch_size = 4;                                   // include/viso.h:25
    c->align_iterations = 100;                                // src/viso.cpp:849
    c->flow_iterations = 10;                                  // src/viso.cpp:266
    c->photometric_error_thresh = (4 * 2) * (4 * 2) * 15 * 15; // include/viso.h:26
}

viso_lk* viso_lk_create(const viso_lk_config* cfg, int32_t* status)
{
    int32_t dummy;
    if (!status) status = &dummy;
    if (!cfg || cfg->struct_size != (int32_t)sizeof(viso_lk_config) || cfg->max_keypoints < 1 || cfg->max_slots < 1
        || cfg->half_patch_size < 1 || 4 * cfg->half_patch_size * cfg->half_patch_size > 32 * kMaxPerLane
        || cfg->align_iterations < 1 || cfg->flow_iterations < 1) {
        *status = VISO_LK_ERR_ARG;
        return nullptr;
    }
    viso_lk* h = new viso_lk();
    h->cfg = *cfg;
    DevImage none;
    memset(&none, 0, sizeof(none));
    h->slots.assign((size_t)cfg->max_slots * kLevels, none);
    h->images.assign((size_t)cfg->max_slots * kLevels, LkBuf());
    *status = lk_init_device(h); // a handle without a device still answers last_error(); compute calls fail
    return h;
}

void viso_lk_destroy(viso_lk* h)
{
    if (!h) return;
    if (h->usable) {
        cudaSetDevice(h->cfg.device);
        cudaStreamSynchronize(h->stream);
        for (auto& b : h->images) cudaFree(b.p);
        cudaFree(h->d_slots.p), cudaFree(h->d_ref_slot.p), cudaFree(h->d_uv_ref.p), cudaFree(h->d_uv_cur.p);
        cudaFree(h->d_kp1.p), cudaFree(h->d_kp2.p), cudaFree(h->d_success.p);
        cudaFree(h->d_pts.p), cudaFree(h->d_src.p), cudaFree(h->d_tgt.p), cudaFree(h->d_fpose.p), cudaFree(h->d_vpose.p), cudaFree(h->d_err.p);
        cudaEventDestroy(h->ev0), cudaEventDestroy(h->ev1);
        cudaStreamDestroy(h->stream);
    }
    delete h;
}

const char* viso_lk_last_error(const viso_lk* h) { return h ? h->err.c_str() : "null handle"; }
int64_t viso_lk_kernel_launches(const viso_lk* h) { return h ? h->launches : 0; }

int32_t viso_lk_set_image(viso_lk* h, int32_t slot, int32_t level, const uint8_t* data, int32_t cols, int32_t rows, int32_t step)
{
    int32_t s = lk_need_device(h);
    if (s) return s;
    if (!data || level < 0 || level >= kLevels || cols < 1 || rows < 1 || step < cols) return lk_fail(h, VISO_LK_ERR_ARG, "set_image: bad argument");
    if (slot < 0 || slot >= h->cfg.max_slots) return lk_fail(h, VISO_LK_ERR_INDEX, "set_image: slot %d outside 0..%d", slot, h->cfg.max_slots - 1);
    if (cols > h->cfg.max_width || rows > h->cfg.max_height)
        return lk_fail(h, VISO_LK_ERR_CAPACITY, "set_image: %dx%d exceeds the handle's %dx%d", cols, rows, h->cfg.max_width, h->cfg.max_height);
    LkBuf& b = h->images[(size_t)slot * kLevels + level];
    const size_t bytes = (size_t)rows * step;
    if ((s = lk_reserve(h, b, bytes))) return s;
    LCU(cudaMemcpyAsync(b.p, data, bytes, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaStreamSynchronize(h->stream)); // the caller may reuse `data` on return
    DevImage& im = h->slots[(size_t)slot * kLevels + level];
    im.data = (const uint8_t*)b.p, im.cols = cols, im.rows = rows, im.step = step, im.pad = 0;
    h->slots_dirty = true;
    return VISO_LK_OK;
}

int32_t viso_lk_set_frame(viso_lk* h, int32_t slot, const uint8_t* data, int32_t cols, int32_t rows, int32_t step)
{
    int32_t s = viso_lk_set_image(h, slot, 0, data, cols, rows, step);
    if (s) return s;
    for (int l = 1; l < kLevels; ++l) { // keyframe.h:40-45
        const DevImage prev = h->slots[(size_t)slot * kLevels + l - 1];
        const int dcols = (int)(prev.cols * 0.5), drows = (int)(prev.rows * 0.5);
        if (dcols < 1 || drows < 1) return lk_fail(h, VISO_LK_ERR_ARG, "set_frame: %dx%d is too small for %d pyramid levels", cols, rows, kLevels);
        LkBuf& b = h->images[(size_t)slot * kLevels + l];
        if ((s = lk_reserve(h, b, (size_t)drows * dcols))) return s;
        h->launches += visoimg::launch_pyr_down(prev.data, prev.cols, prev.rows, prev.step, (uint8_t*)b.p, dcols, drows, dcols, h->stream);
        LCU(cudaGetLastError());
        DevImage& im = h->slots[(size_t)slot * kLevels + l];
        im.data = (const uint8_t*)b.p, im.cols = dcols, im.rows = drows, im.step = dcols, im.pad = 0;
    }
    h->slots_dirty = true;
    LCU(cudaStreamSynchronize(h->stream));
    return VISO_LK_OK;
}

int32_t viso_lk_get_image(viso_lk* h, int32_t slot, int32_t level, uint8_t* out, int64_t capacity, int32_t* cols, int32_t* rows)
{
    int32_t s = lk_need_device(h);
    if (s) return s;
    if (!out || !cols || !rows || level < 0 || level >= kLevels) return lk_fail(h, VISO_LK_ERR_ARG, "get_image: bad argument");
    if ((s = lk_check_slot(h, slot, level, level))) return s;
    const DevImage& im = h->slots[(size_t)slot * kLevels + level];
    if (capacity < (int64_t)im.cols * im.rows) return lk_fail(h, VISO_LK_ERR_CAPACITY, "get_image: %dx%d does not fit %lld bytes", im.cols, im.rows, (long long)capacity);
    LCU(cudaMemcpy2DAsync(out, im.cols, im.data, im.step, im.cols, im.rows, cudaMemcpyDeviceToHost, h->stream));
    LCU(cudaStreamSynchronize(h->stream));
    *cols = im.cols, *rows = im.rows;
    return VISO_LK_OK;
}

int32_t viso_lk_align(viso_lk* h, int32_t n, const int32_t* ref_slot, int32_t cur_slot, const double* uv_ref, double* uv_cur,
                      uint8_t* success, float* device_ms)
{
    return lk_align_run(h, kLevels - 1, 0, n, ref_slot, cur_slot, uv_ref, uv_cur, success, device_ms);
}

int32_t viso_lk_align_level(viso_lk* h, int32_t level, int32_t n, const int32_t* ref_slot, int32_t cur_slot, const double* uv_ref,
                            double* uv_cur, uint8_t* success)
{
    if (level < 0 || level >= kLevels) return lk_fail(h, VISO_LK_ERR_ARG, "align_level: level %d", level);
    return lk_align_run(h, level, level, n, ref_slot, cur_slot, uv_ref, uv_cur, success, nullptr);
}

int32_t viso_lk_optical_flow(viso_lk* h, int32_t slot1, int32_t slot2, int32_t n, const float* kp1, float* kp2,
                             int32_t have_initial, int32_t inverse, uint8_t* success, float* device_ms)
{
    return lk_flow_run(h, kLevels - 1, 0, 1, slot1, slot2, n, kp1, kp2, have_initial, inverse, success, device_ms);
}

int32_t viso_lk_optical_flow_level(viso_lk* h, int32_t level, int32_t slot1, int32_t slot2, int32_t n, const float* kp1,
                                   float* kp2, int32_t have_initial, int32_t inverse, uint8_t* success)
{
    if (level < 0 || level >= kLevels) return lk_fail(h, VISO_LK_ERR_ARG, "optical_flow_level: level %d", level);
    return lk_flow_run(h, level, level, 0, slot1, slot2, n, kp1, kp2, have_initial, inverse, success, nullptr);
}

int32_t viso_lk_edge_direct_projection(viso_lk* h, int32_t n, const double* points, const int32_t* src_slot, const int32_t* tgt_slot,
                                       const double* frame_poses, const double* vertex_poses, double fx, double fy, double cx, double cy,
                                       double* errors, uint8_t* level, float* device_ms)
{
    int32_t s = lk_need_device(h);
    if (s) return s;
    if (n < 0 || (n > 0 && (!points || !src_slot || !tgt_slot || !frame_poses || !vertex_poses || !errors || !level)))
        return lk_fail(h, VISO_LK_ERR_ARG, "edge_direct_projection: bad argument");
    if (n > h->cfg.max_keypoints) return lk_fail(h, VISO_LK_ERR_CAPACITY, "edge_direct_projection: %d edges exceed the pool of %d", n, h->cfg.max_keypoints);
    if (device_ms) *device_ms = 0.f;
    if (n == 0) return VISO_LK_OK;
    for (int32_t e = 0; e < n; ++e)
        if ((s = lk_check_slot(h, src_slot[e], 0, 0)) || (s = lk_check_slot(h, tgt_slot[e], 0, 0))) return s;
    const size_t ns = (size_t)h->cfg.max_slots, cap = (size_t)h->cfg.max_keypoints;
    if (!lk_finite(points, 3 * (size_t)n) || !lk_finite(frame_poses, 7 * ns) || !lk_finite(vertex_poses, 7 * ns))
        return lk_fail(h, VISO_LK_ERR_NONFINITE, "edge_direct_projection: NaN/Inf in the points or poses");
    if ((s = lk_sync_slots(h))) return s;
    if ((s = lk_reserve(h, h->d_pts, sizeof(double) * 3 * cap)) || (s = lk_reserve(h, h->d_src, sizeof(int32_t) * cap))
        || (s = lk_reserve(h, h->d_tgt, sizeof(int32_t) * cap)) || (s = lk_reserve(h, h->d_fpose, sizeof(double) * 7 * ns))
        || (s = lk_reserve(h, h->d_vpose, sizeof(double) * 7 * ns)) || (s = lk_reserve(h, h->d_err, sizeof(double) * 16 * cap))
        || (s = lk_reserve(h, h->d_success, cap)))
        return s;
    LCU(cudaMemcpyAsync(h->d_pts.p, points, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaMemcpyAsync(h->d_src.p, src_slot, sizeof(int32_t) * n, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaMemcpyAsync(h->d_tgt.p, tgt_slot, sizeof(int32_t) * n, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaMemcpyAsync(h->d_fpose.p, frame_poses, sizeof(double) * 7 * ns, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaMemcpyAsync(h->d_vpose.p, vertex_poses, sizeof(double) * 7 * ns, cudaMemcpyHostToDevice, h->stream));
    LCU(cudaEventRecord(h->ev0, h->stream));
    const int grid = (int)(((size_t)n * 16 + kThreads - 1) / kThreads);
    k_edge_direct_projection<<<grid, kThreads, 0, h->stream>>>((const DevImage*)h->d_slots.p, n, (const double*)h->d_pts.p,
        (const int32_t*)h->d_src.p, (const int32_t*)h->d_tgt.p, (const double*)h->d_fpose.p, (const double*)h->d_vpose.p, fx, fy, cx, cy,
        (double*)h->d_err.p, (uint8_t*)h->d_success.p);
    LCU(cudaGetLastError());
    ++h->launches;
    LCU(cudaEventRecord(h->ev1, h->stream));
    LCU(cudaMemcpyAsync(errors, h->d_err.p, sizeof(double) * 16 * n, cudaMemcpyDeviceToHost, h->stream));
    LCU(cudaMemcpyAsync(level, h->d_success.p, n, cudaMemcpyDeviceToHost, h->stream));
    LCU(cudaStreamSynchronize(h->stream));
    if (device_ms) LCU(cudaEventElapsedTime(device_ms, h->ev0, h->ev1));
    return VISO_LK_OK;
}

} // extern "C"
