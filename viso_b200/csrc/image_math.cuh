// image_math.cuh — device-side image sampling shared by the direct pose estimator (direct_pose.cu)
// and the Lucas-Kanade patch alignment (lk_align.cu), statement by statement as the reference
// writes it (citations relative to /root/reference):
//   GetPixelValue   include/keyframe.h:50-57 and include/common.h:35-42 (identical arithmetic)
//   GetGradient     include/keyframe.h:59-64 and include/common.h:44-49
//   IsInside        include/keyframe.h:76-79
#ifndef VISO_IMAGE_MATH_CUH
#define VISO_IMAGE_MATH_CUH

#include <cuda_runtime.h>
#include <stdint.h>

namespace visoimg {

struct DevImage { // one pyramid level of one frame (cv::Mat CV_8UC1), resident in HBM
    const uint8_t* data;
    int32_t cols, rows, step, pad;
};

// bilinear sample: int() truncation for the address, floor() for the weights.  Reads that leave
// the cv::Mat buffer in the reference (undefined behaviour there) are clamped into it.
__device__ __forceinline__ double pixel_value(const DevImage& im, double x, double y)
{
    const long long base = (long long)(int)y * im.step + (long long)(int)x;
    const long long last = (long long)im.rows * im.step - 1;
    const double xx = x - floor(x);
    const double yy = y - floor(y);
    const long long i0 = min(max(base, 0LL), last), i1 = min(max(base + 1, 0LL), last);
    const long long i2 = min(max(base + im.step, 0LL), last), i3 = min(max(base + im.step + 1, 0LL), last);
    const double d0 = (double)__ldg(im.data + i0), d1 = (double)__ldg(im.data + i1);
    const double d2 = (double)__ldg(im.data + i2), d3 = (double)__ldg(im.data + i3);
    return (1 - xx) * (1 - yy) * d0 + xx * (1 - yy) * d1 + (1 - xx) * yy * d2 + xx * yy * d3;
}

__device__ __forceinline__ void gradient(const DevImage& im, double u, double v, double& gx, double& gy)
{
    gx = 0.5 * (pixel_value(im, u + 1, v) - pixel_value(im, u - 1, v));
    gy = 0.5 * (pixel_value(im, u, v + 1) - pixel_value(im, u, v - 1));
}

__device__ __forceinline__ bool is_inside(const DevImage& im, double u, double v)
{
    return u >= 0 && u < im.cols && v >= 0 && v < im.rows;
}

// pyramid.cu: one cv::pyrDown step (CV_8UC1, Size(cols * 0.5, rows * 0.5)) on the device; returns the launches made
int launch_pyr_down(const uint8_t* src, int cols, int rows, int step, uint8_t* dst, int dcols, int drows, int dstep, cudaStream_t st);

} // namespace visoimg
#endif
