// tools/chol_bench.cu — stand-alone timing of the dense reduced-system solver (ba_dense.cu) on a
// random SPD matrix, with CTA 0's time stamps at the phase boundaries.  Not part of the product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DVISO_CHOL_PROFILE -I viso_b200/csrc \
//        tools/chol_bench.cu -o tools/_build/chol_bench && tools/_build/chol_bench 1194
#include "../viso_b200/csrc/ba_dense.cu"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace visoba;

int main(int argc, char** argv)
{
    const int n = argc > 1 ? atoi(argv[1]) : 1194;
    const int reps = argc > 2 ? atoi(argv[2]) : 10;
    const int band = argc > 3 ? atoi(argv[3]) : n; // half bandwidth in columns (entries with i - j > band are zero)
    const int ld = (n + 8) & ~7;
    std::vector<double> A((size_t)(n + 1) * ld, 0.0), M((size_t)n * 8);
    unsigned long long s = 88172645463325252ull;
    auto rnd = [&] { s ^= s << 13, s ^= s >> 7, s ^= s << 17; return (double)(s >> 11) / 9007199254740992.0 - 0.5; };
    for (auto& v : M) v = rnd();
    for (int i = 0; i < n; ++i)
        for (int j = 0; j <= i; ++j) {
            double a = i == j ? 4.0 : 0.0;
            for (int k = 0; k < 8; ++k) a += M[(size_t)i * 8 + k] * M[(size_t)j * 8 + k];
            if (i - j > band) a = 0.0; else if (i != j) a *= 0.1; // diagonally dominant: SPD with any band
            A[(size_t)i * ld + j] = a;
        }
    std::vector<double> xt(n), b(n, 0.0);
    for (auto& v : xt) v = rnd();
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) b[i] += (j <= i ? A[(size_t)i * ld + j] : A[(size_t)j * ld + i]) * xt[j];
    for (int j = 0; j < n; ++j) A[(size_t)n * ld + j] = b[j];

    LargeSolve L;
    L.n = n, L.ld = ld;
    double* dA0;
    cudaMalloc(&dA0, A.size() * 8);
    cudaMalloc(&L.A, A.size() * 8);
    cudaMalloc(&L.Dd, large_solve_diag_doubles(n) * 8);
    cudaMalloc(&L.x, n * 8);
    cudaMalloc(&L.fail, 4);
    const int nb64 = (n + 63) / 64;
    std::vector<int> first(nb64), last(nb64);
    for (int i = 0; i < nb64; ++i) first[i] = std::max(0, (64 * i - band) / 64);
    for (int i = nb64 - 2; i >= 0; --i) first[i] = std::min(first[i], first[i + 1]);
    for (int k = 0, i = 0; k < nb64; ++k) {
        while (i + 1 < nb64 && first[i + 1] <= k) ++i;
        last[k] = std::max(i, k);
    }
    int *dfirst, *dlast;
    cudaMalloc(&dfirst, nb64 * 4), cudaMalloc(&dlast, nb64 * 4);
    cudaMemcpy(dfirst, first.data(), nb64 * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dlast, last.data(), nb64 * 4, cudaMemcpyHostToDevice);
    L.first64 = dfirst, L.last64 = dlast;
    cudaMemset(L.fail, 0, 4);
    cudaMemcpy(dA0, A.data(), A.size() * 8, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaMemcpy(L.A, dA0, A.size() * 8, cudaMemcpyDeviceToDevice);
        cudaEventRecord(e0);
        if (!launch_chol_coop(L, 0)) { printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
        cudaEventRecord(e1);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    std::vector<double> x(n);
    int fail = 0;
    cudaMemcpy(x.data(), L.x, n * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(&fail, L.fail, 4, cudaMemcpyDeviceToHost);
    double err = 0.0;
    for (int i = 0; i < n; ++i) err = fmax(err, fabs(x[i] - xt[i]));
    printf("n %d  best %.1f us  fail %d  max |x - x_true| %.3e\n", n, best * 1e3, fail, err);
#ifdef VISO_CHOL_PROFILE
    const int steps = (n + 63) / 64;
    std::vector<unsigned long long> p(16 * 512);
    cudaMemcpyFromSymbol(p.data(), g_chol_prof, sizeof(unsigned long long) * 16 * 512);
    const char* names[9] = {"load", "chol11", "L21", "A22", "chol22", "panel", "sync1", "update", "sync2"};
    double tot[9] = {};
    for (int k = 0; k < steps && k < 511; ++k) {
        const unsigned long long* q = &p[16 * k];
        for (int i = 0; i < 9; ++i) tot[i] += (double)(q[i + 1] - q[i]) * 1e-3;
        if (k == 1) {
            printf("step 1:");
            for (int i = 0; i < 9; ++i) printf(" %s %.1f", names[i], (double)(q[i + 1] - q[i]) * 1e-3);
            printf(" us\n");
        }
    }
    printf("totals:");
    for (int i = 0; i < 9; ++i) printf(" %s %.1f", names[i], tot[i]);
    double sum = 0;
    for (int i = 0; i < 9; ++i) sum += tot[i];
    printf(" | backsub+rest %.1f us\n", best * 1e3 - sum);
#endif
    return err < 1e-8 && !fail ? 0 : 1;
}
