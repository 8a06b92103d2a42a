// tools/chol_bench.cu — stand-alone check and timing of the large reduced-system solver (ba_dense.cu) on a random
// block-banded SPD system given as packed 6x6 S blocks, exactly as the adjuster hands it over: plan (ba_solve_plan.h) ->
// zero + assemble -> the cooperative kernel -> x, against the known solution.  Not part of the product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 [-DVISO_CHOL_PROFILE] -I viso_b200/csrc \
//        tools/chol_bench.cu -o tools/_build/chol_bench
//   chol_bench <free poses> <span> <max segments> <force> [reps]      e.g.  199 199 1 0   (C3)    1999 25 8 0   (C4)
#include "../viso_b200/csrc/ba_dense.cu"
#include "../viso_b200/csrc/ba_solve_plan.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace visoba;

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            printf("%s: %s (line %d)\n", #x, cudaGetErrorString(e_), __LINE__);                \
            return 2;                                                                          \
        }                                                                                      \
    } while (0)

int main(int argc, char** argv)
{
    const int nf = argc > 1 ? atoi(argv[1]) : 199, span = argc > 2 ? atoi(argv[2]) : nf;
    const int max_seg = argc > 3 ? atoi(argv[3]) : 8, force = argc > 4 ? atoi(argv[4]) : 0, reps = argc > 5 ? atoi(argv[5]) : 10;
    const int n = 6 * nf;
    unsigned long long s = 88172645463325252ull + (unsigned long long)nf * 7919u + (unsigned long long)span;
    auto rnd = [&] { s ^= s << 13, s ^= s >> 7, s ^= s << 17; return (double)(s >> 11) / 9007199254740992.0 - 0.5; };
    std::vector<int32_t> blocks;
    std::vector<double> Sblk;
    for (int I = 0; I < nf; ++I)
        for (int J = std::max(0, I - span); J <= I; ++J) {
            blocks.push_back(I), blocks.push_back(J);
            for (int r = 0; r < 6; ++r)
                for (int c = 0; c < 6; ++c) {
                    double v = 0.3 * rnd();
                    if (I == J) v = r == c ? 6.0 * (std::min(span, nf) + 2) : (c < r ? v : Sblk[Sblk.size() - (size_t)(6 * r + c) + (size_t)(6 * c + r)]);
                    Sblk.push_back(v);
                }
        }
    const size_t nblk = blocks.size() / 2;
    std::vector<double> xt((size_t)n), bvec((size_t)n, 0.0);
    for (auto& v : xt) v = rnd();
    for (size_t k = 0; k < nblk; ++k) { // b = S x_true from the packed blocks
        const int I = blocks[2 * k], J = blocks[2 * k + 1];
        for (int r = 0; r < 6; ++r)
            for (int c = 0; c < 6; ++c) {
                const double v = Sblk[36 * k + (size_t)(6 * r + c)];
                if (I == J) {
                    if (c <= r) {
                        bvec[(size_t)(6 * I + r)] += v * xt[(size_t)(6 * J + c)];
                        if (c < r) bvec[(size_t)(6 * J + c)] += v * xt[(size_t)(6 * I + r)];
                    }
                } else {
                    bvec[(size_t)(6 * I + r)] += v * xt[(size_t)(6 * J + c)];
                    bvec[(size_t)(6 * J + c)] += v * xt[(size_t)(6 * I + r)];
                }
            }
    }
    CK(cudaSetDevice(0));
    const SolvePlan plan = make_solve_plan(nf, blocks, large_solve_cta_count(), max_seg, force != 0);
    const size_t nt = plan.tasks.size();
    // device memory
    size_t nd = 0, ni = 0, ns = 0;
    for (const PlanTask& t : plan.tasks) nd += t.a_doubles() + t.dd_doubles() + (size_t)t.nrows, ni += 2 * (size_t)(t.nrows / 64), ns += (size_t)(t.n_elim / 64);
    ni += 3 * (size_t)nf + nt + 1;
    double *dD, *dx, *dS, *dbs;
    int32_t *dI, *dfail, *dfree;
    int4* dsteps;
    int2* dij;
    CholTask* dtasks;
    unsigned int* dbars;
    CK(cudaMalloc(&dD, nd * 8));
    CK(cudaMalloc(&dx, (size_t)n * 8));
    CK(cudaMalloc(&dI, ni * 4));
    CK(cudaMalloc(&dsteps, ns * 16));
    CK(cudaMalloc(&dtasks, nt * sizeof(CholTask)));
    CK(cudaMalloc(&dbars, nt * 4 * kBarsPerTask));
    CK(cudaMalloc(&dfail, 4));
    CK(cudaMalloc(&dS, Sblk.size() * 8));
    CK(cudaMalloc(&dbs, (size_t)n * 8));
    CK(cudaMalloc(&dij, nblk * 8));
    CK(cudaMalloc(&dfree, (size_t)nf * 4));
    std::vector<CholTask> tasks(nt);
    std::vector<int32_t> ints, steps;
    {
        double* pd = dD;
        int4* ps = dsteps;
        size_t io = 0;
        for (size_t t = 0; t < nt; ++t) {
            const PlanTask& q = plan.tasks[t];
            CholTask& c = tasks[t];
            memset(&c, 0, sizeof(c));
            c.A = pd, pd += q.a_doubles(), c.Dd = pd, pd += q.dd_doubles(), c.x = pd, pd += (size_t)q.nrows;
            const size_t nb = (size_t)(q.nrows / 64);
            c.flags = dI + io, io += nb, ints.insert(ints.end(), nb, 0);
            c.first64 = dI + io, io += nb, ints.insert(ints.end(), q.first64.begin(), q.first64.end());
            c.steps = ps, ps += q.n_elim / 64, steps.insert(steps.end(), q.steps.begin(), q.steps.end());
            c.ld = q.ld, c.n_elim = q.n_elim, c.nrows = q.nrows, c.n_real = q.n_real, c.cta0 = q.cta0, c.ncta = q.ncta, c.sL = q.sL, c.sR = q.sR;
        }
        LargeSolve L;
        memset(&L, 0, sizeof(L));
        L.n = n, L.n_seg = plan.n_seg, L.sep_cols = plan.sep_cols, L.total_rows = plan.total_rows, L.tasks = dtasks;
        L.cam_task = dI + io, io += (size_t)nf, ints.insert(ints.end(), plan.cam_task.begin(), plan.cam_task.end());
        L.cam_off = dI + io, io += (size_t)nf, ints.insert(ints.end(), plan.cam_off.begin(), plan.cam_off.end());
        L.cam_sep = dI + io, io += (size_t)nf, ints.insert(ints.end(), plan.cam_sep.begin(), plan.cam_sep.end());
        L.row_ptr = dI + io, io += nt + 1, ints.insert(ints.end(), plan.row_ptr.begin(), plan.row_ptr.end());
        L.x = dx, L.fail = dfail, L.bars = dbars;
        CK(cudaMemcpy(dI, ints.data(), ints.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dsteps, steps.data(), steps.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dtasks, tasks.data(), nt * sizeof(CholTask), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dS, Sblk.data(), Sblk.size() * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dbs, bvec.data(), (size_t)n * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dij, blocks.data(), nblk * 8, cudaMemcpyHostToDevice));
        std::vector<int32_t> fr((size_t)nf);
        for (int f = 0; f < nf; ++f) fr[(size_t)f] = f;
        CK(cudaMemcpy(dfree, fr.data(), (size_t)nf * 4, cudaMemcpyHostToDevice));
        DevProblem P;
        memset(&P, 0, sizeof(P));
        P.n_blocks = (int32_t)nblk, P.n_cams = nf, P.blk_ij = dij, P.S_blk = dS, P.bs = dbs, P.cam_free = dfree;
        cudaEvent_t e0, e1, e2;
        cudaEventCreate(&e0), cudaEventCreate(&e1), cudaEventCreate(&e2);
        float best = 1e30f, best_asm = 1e30f;
        const long long warps = (long long)P.n_blocks + P.n_cams;
        for (int r = 0; r < reps; ++r) {
            cudaEventRecord(e0);
            k_dense_zero<<<L.total_rows, 256>>>(L);
            k_dense_assemble<<<(unsigned)((warps * 32 + 255) / 256), 256>>>(P, L);
            cudaEventRecord(e1);
            if (!launch_chol_coop(L, 0)) return printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())), 1;
            cudaEventRecord(e2);
            CK(cudaDeviceSynchronize());
            float ms, ms2;
            cudaEventElapsedTime(&ms2, e0, e1), cudaEventElapsedTime(&ms, e1, e2);
            best = std::min(best, ms), best_asm = std::min(best_asm, ms2);
        }
        int fail = 0;
        CK(cudaMemcpy(&fail, dfail, 4, cudaMemcpyDeviceToHost));
        double err = 0.0;
        for (size_t t = 0; t < nt; ++t) {
            std::vector<double> x((size_t)plan.tasks[t].nrows);
            CK(cudaMemcpy(x.data(), tasks[t].x, x.size() * 8, cudaMemcpyDeviceToHost));
            for (int f = 0; f < nf; ++f)
                if (plan.cam_task[(size_t)f] == (int)t)
                    for (int r2 = 0; r2 < 6; ++r2) {
                        const double e = fabs(x[(size_t)(plan.cam_off[(size_t)f] + r2)] - xt[(size_t)(6 * f + r2)]);
                        if (!(e <= err)) err = e;
                    }
        }
        printf("nf %d n %d span %d segments %d (separators of %d poses) steps:", nf, n, plan.span, plan.n_seg, plan.sep_cams);
        for (const PlanTask& T : plan.tasks) printf(" %d", T.n_elim / 64);
        printf("  |  zero+assemble %.1f us  solve %.1f us  fail %d  max |x - x_true| %.3e\n", best_asm * 1e3, best * 1e3, fail, err);
#ifdef VISO_CHOL_PROFILE
        const int nsteps = plan.tasks.back().n_elim / 64;
        std::vector<unsigned long long> p(16 * 512);
        cudaMemcpyFromSymbol(p.data(), g_chol_prof, sizeof(unsigned long long) * 16 * 512);
        long long c64[32];
        cudaMemcpyFromSymbol(c64, g_c64_prof, sizeof(c64));
        printf("cta_chol64 (CTA 0, last call), cycles: first 8x8 %lld, transpose %lld; per sub-step panel / next-diag+chol8 / wait:", c64[28], c64[29]);
        for (int i = 0; i < 7; ++i) printf("  %lld/%lld/%lld", c64[4 * i], c64[4 * i + 1], c64[4 * i + 2]);
        printf("\n  warp 0 after its panel rows, last sub-step: arrive %lld, next-diagonal update %lld cycles\n", c64[30], c64[31]);
        // lead (CTA 0): marks 0..5 = step start, diagonal entries loaded (after U of the previous step), panel there (B),
        // next diagonal block formed, factorised, published.  Worker (CTA 1): marks 8..13 = step start, factor loaded,
        // panel done, B passed, update done, U passed.
        const char* ln[5] = { "wait U + own entries", "wait B", "tile", "chol64", "publish" };
        const char* wn[5] = { "wait F + load", "panel", "wait B", "update", "wait U" };
        double lt[5] = {}, wt[5] = {};
        for (int k = 0; k + 1 < nsteps && k < 511; ++k) {
            const unsigned long long* q = &p[16 * k];
            for (int i = 0; i < 5; ++i) lt[i] += (double)(q[i + 1] - q[i]) * 1e-3, wt[i] += (double)(q[9 + i] - q[8 + i]) * 1e-3;
        }
        printf("T lead CTA, totals over %d steps (us):", nsteps - 1);
        for (int i = 0; i < 5; ++i) printf("  %s %.1f", ln[i], lt[i]);
        printf("\nT worker CTA 1, totals (us):");
        for (int i = 0; i < 5; ++i) printf("  %s %.1f", wn[i], wt[i]);
        if (nsteps < 511) printf("  | backward %.1f us", (double)(p[16 * nsteps + 1] - p[16 * nsteps]) * 1e-3);
        printf("\n");
#endif
        return err < 1e-8 && !fail ? 0 : 1;
    }
}
