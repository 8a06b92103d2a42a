// tools/lat.cu — dependent-issue latencies on the device (cycles): DFMA, DMUL, rsqrt(double), 1/x, sqrt, SHFL of a double.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP> __global__ void k(double* out, long long* cyc, double seed, int n)
{
    double x = seed + threadIdx.x * 1e-9, y = 1.0000001;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        if (OP == 0) x = fma(x, y, 1e-9);
        if (OP == 1) x = x * y;
        if (OP == 2) x = rsqrt(x) + 1.0;
        if (OP == 3) x = 1.0 / x + 0.5;
        if (OP == 4) x = sqrt(x) + 1.0;
        if (OP == 5) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31);
        if (OP == 6) x = x + y;
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) *cyc = t1 - t0;
    out[threadIdx.x] = x;
}
int main()
{
    double* out; long long* cyc; cudaMalloc(&out, 256); cudaMalloc(&cyc, 8);
    const char* names[] = {"DFMA", "DMUL", "rsqrt+add", "div+add", "sqrt+add", "SHFL.f64", "DADD"};
    const int n = 4096;
    long long c;
#define RUN(OP) k<OP><<<1, 32>>>(out, cyc, 1.5, n); k<OP><<<1, 32>>>(out, cyc, 1.5, n); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); printf("%-10s %.1f cycles\n", names[OP], (double)c / n);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6)
    return 0;
}
