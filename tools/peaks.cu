// peaks.cu — measures the CUDA-core FMA roofs the bundle-adjustment kernels are bounded by:
// fp64 DFMA and fp32 FFMA throughput of the whole chip (no tensor cores), plus a plain
// device-to-device copy for reference.  SURVEY H1: the explicit Schur complement is above the
// fp64 ridge, so bench.py reports the fraction of this roof next to the HBM fraction.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/peaks tools/peaks.cu
//   tools/_build/peaks            -> one JSON line on stdout
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

template <typename T, int CHAINS>
__global__ void __launch_bounds__(256) k_fma(T* out, int iters, T a, T b)
{
    T acc[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) acc[i] = (T)(threadIdx.x + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) acc[i] = acc[i] * a + b;
    }
    T s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += acc[i];
    if (s == (T)123456789) out[0] = s; // never true; keeps the chains alive
}

template <typename T> static double fma_tflops(int sms)
{
    constexpr int CH = 16;
    T* d = nullptr;
    cudaMalloc(&d, sizeof(T));
    const int iters = 4096, blocks = sms * 8, threads = 256;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        k_fma<T, CH><<<blocks, threads>>>(d, iters, (T)1.0000001, (T)1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * (double)blocks * threads * CH * (double)iters;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaFree(d);
    return best;
}

static double copy_gbs()
{
    const size_t n = (size_t)1 << 30;
    char *a = nullptr, *b = nullptr;
    if (cudaMalloc(&a, n) != cudaSuccess || cudaMalloc(&b, n) != cudaSuccess) return 0.0;
    cudaMemset(a, 1, n);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        cudaMemcpyAsync(b, a, n, cudaMemcpyDeviceToDevice);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double gbs = 2.0 * (double)n / (ms * 1e-3) / 1e9;
        if (rep > 0 && gbs > best) best = gbs;
    }
    cudaFree(a);
    cudaFree(b);
    return best;
}

int main()
{
    int dev = 0;
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) {
        fprintf(stderr, "no CUDA device\n");
        return 1;
    }
    const double f64 = fma_tflops<double>(p.multiProcessorCount);
    const double f32 = fma_tflops<float>(p.multiProcessorCount);
    const double cp = copy_gbs();
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_khz\": %d, \"fp64_fma_tflops\": %.3f, \"fp32_fma_tflops\": %.3f, "
           "\"d2d_copy_gbs\": %.1f, \"how\": \"16 independent FMA chains per thread, 8 CTAs x 256 threads per SM, best of 5; "
           "cudaMemcpy D2D 1 GiB read+write, best of 5\"}\n",
        p.name, p.multiProcessorCount, clk, f64, f32, cp);
    return 0;
}
