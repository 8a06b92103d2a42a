// pyramid.cu — the image pyramid of a Keyframe, built on the device.
//
// Reference: Keyframe's constructor (include/keyframe.h:29-45) keeps the frame as level 0 and calls
// cv::pyrDown(pyramids_[i-1], pyr, Size(cols * pyramid_scale_, rows * pyramid_scale_)) three times
// (nr_pyramids_ = 4, pyramid_scale_ = 0.5, keyframe.h:19-20).  OpenCV is a dependency that is not under
// /root/reference; its published algorithm for CV_8UC1 (imgproc/pyramids.cpp) is restated here:
// separable 5-tap kernel [1 4 6 4 1] centred on (2x, 2y), BORDER_REFLECT_101, integer sums,
// dst = (sum + 128) >> 8.  Integer work: bit-exact against cv2.pyrDown's own outputs
// (tests/golden/pyrdown_cv2.npz).
//
// HBM-bound and tiny (a 640x480 level reads 300 kB, writes 75 kB): one thread per output pixel,
// a warp writes 32 consecutive bytes and reads a 68-byte span of five source rows through L1.
#include "image_math.cuh"

namespace visoimg {

namespace {

__device__ __forceinline__ int reflect101(int i, int n)
{
    i = i < 0 ? -i : i;
    return i >= n ? 2 * (n - 1) - i : i;
}

__global__ void __launch_bounds__(256) k_pyr_down(const uint8_t* __restrict__ src, int cols, int rows, int step,
    uint8_t* __restrict__ dst, int dcols, int drows, int dstep)
{
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= dcols || y >= drows) return;
    int xs[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) xs[j] = reflect101(2 * x + j - 2, cols);
    const int w[5] = { 1, 4, 6, 4, 1 };
    int sum = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const uint8_t* row = src + (size_t)reflect101(2 * y + i - 2, rows) * step;
        const int h = (int)__ldg(row + xs[0]) + (int)__ldg(row + xs[4]) + 4 * ((int)__ldg(row + xs[1]) + (int)__ldg(row + xs[3]))
            + 6 * (int)__ldg(row + xs[2]);
        sum += w[i] * h;
    }
    dst[(size_t)y * dstep + x] = (uint8_t)((sum + 128) >> 8);
}

} // namespace

int launch_pyr_down(const uint8_t* src, int cols, int rows, int step, uint8_t* dst, int dcols, int drows, int dstep, cudaStream_t st)
{
    if (dcols < 1 || drows < 1) return 0;
    dim3 grid((dcols + 31) / 32, (drows + 7) / 8);
    k_pyr_down<<<grid, 256, 0, st>>>(src, cols, rows, step, dst, dcols, drows, dstep);
    return 1;
}

} // namespace visoimg
