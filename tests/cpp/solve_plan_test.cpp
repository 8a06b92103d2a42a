// solve_plan_test.cpp — CPU check of the large solver's PLAN (viso_b200/csrc/ba_solve_plan.h): a sequential emulation of
// what ba_dense.cu does with the plan (assemble the task matrices from packed S blocks, eliminate every segment over the
// row ranges its steps list, add the border blocks into the separator system, solve it, back-substitute the interiors),
// compared with a plain dense Cholesky solve of the same block-banded SPD system.  It follows the kernel loop for loop
// (64-column steps, two row ranges per step, the right-hand side as one more row), so an index error in the plan or in
// the scheme shows up here, on the CPU, before any GPU time is spent.  Test infrastructure only.
//   g++ -O2 -std=c++17 -I viso_b200/csrc tests/cpp/solve_plan_test.cpp -o tests/cpp/_build/solve_plan_test
//   solve_plan_test <free poses> <span> <max segments> <force 0|1> <ctas>
#include "ba_solve_plan.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

using namespace visoba;

struct TaskMem {
    std::vector<double> A, x;
    std::vector<int> flags;
};

static bool chol64(double* D) // 64x64 lower, row-major stride 64, in place
{
    for (int j = 0; j < 64; ++j) {
        double d = D[j * 64 + j];
        for (int k = 0; k < j; ++k) d -= D[j * 64 + k] * D[j * 64 + k];
        if (!(d > 0.0)) return false;
        const double l = std::sqrt(d);
        D[j * 64 + j] = l;
        for (int i = j + 1; i < 64; ++i) {
            double s = D[i * 64 + j];
            for (int k = 0; k < j; ++k) s -= D[i * 64 + k] * D[j * 64 + k];
            D[i * 64 + j] = s / l;
        }
    }
    return true;
}

static int step_row(int q, int base, int n1, int r2b, int n2, int rhs) { return q < n1 ? base + q : (q < n1 + n2 ? r2b + (q - n1) : rhs); }

static bool task_factor(const PlanTask& T, TaskMem& M, std::vector<double>& Dd)
{
    const int ld = T.ld, nrows = T.nrows;
    double* A = M.A.data();
    Dd.assign((size_t)(T.n_elim / 64) * 4096, 0.0);
    for (int k = 0; k < T.n_elim / 64; ++k) {
        const int j0 = 64 * k, base = j0 + 64;
        const int r1e = T.steps[4 * k], r2b = T.steps[4 * k + 1], r2e = T.steps[4 * k + 2];
        const int n1 = std::max(r1e - base, 0), n2 = std::max(r2e - r2b, 0);
        double D[4096];
        for (int r = 0; r < 64; ++r)
            for (int c = 0; c < 64; ++c) D[r * 64 + c] = c <= r ? A[(size_t)(j0 + r) * ld + j0 + c] : 0.0;
        if (!chol64(D)) return false;
        for (int i = 0; i < 4096; ++i) Dd[(size_t)k * 4096 + i] = D[i];
        const int nrow = n1 + n2 + 1;
        for (int q = 0; q < nrow; ++q) { // panel: x L^T = a
            double* a = A + (size_t)step_row(q, base, n1, r2b, n2, nrows) * ld + j0;
            for (int c = 0; c < 64; ++c) {
                double s = a[c];
                for (int t = 0; t < c; ++t) s -= a[t] * D[c * 64 + t];
                a[c] = s / D[c * 64 + c];
            }
        }
        const int n1t = n1 / 64, nt = n1t + n2 / 64;
        for (int ti = 0; ti <= nt; ++ti) // ti == nt: the right-hand side row
            for (int tl = 0; tl < nt && (ti == nt || tl <= ti); ++tl) {
                const bool rhs = ti == nt;
                const int i0 = rhs ? nrows : (ti < n1t ? base + 64 * ti : r2b + 64 * (ti - n1t));
                const int l0 = tl < n1t ? base + 64 * tl : r2b + 64 * (tl - n1t);
                for (int r = 0; r < (rhs ? 1 : 64); ++r)
                    for (int c = 0; c < 64; ++c) {
                        if (!rhs && ti == tl && c > r) continue;
                        double s = 0.0;
                        for (int kk = 0; kk < 64; ++kk) s += A[(size_t)(i0 + r) * ld + j0 + kk] * A[(size_t)(l0 + c) * ld + j0 + kk];
                        A[(size_t)(i0 + r) * ld + l0 + c] -= s;
                    }
            }
    }
    return true;
}

static bool task_backward(const PlanTask& T, TaskMem& M, const std::vector<double>& Dd)
{
    const int ld = T.ld, nrows = T.nrows;
    const double* A = M.A.data();
    for (int b = T.n_elim / 64 - 1; b >= 0; --b) {
        const int cb = 64 * b;
        const int r1e = T.steps[4 * b], r2b = T.steps[4 * b + 1], r2e = T.steps[4 * b + 2];
        double y[64];
        for (int c = 0; c < 64; ++c) y[c] = A[(size_t)nrows * ld + cb + c];
        const int nb1 = std::max(r1e - (cb + 64), 0) / 64, nb2 = std::max(r2e - r2b, 0) / 64;
        for (int i = nb1 + nb2 - 1; i >= 0; --i) {
            const int rb = i >= nb1 ? r2b + 64 * (i - nb1) : cb + 64 + 64 * i;
            if (!M.flags[rb / 64]) return false; // would deadlock on the device
            for (int r = 0; r < 64; ++r)
                for (int c = 0; c < 64; ++c) y[c] -= A[(size_t)(rb + r) * ld + cb + c] * M.x[rb + r];
        }
        const double* D = &Dd[(size_t)b * 4096];
        for (int j = 63; j >= 0; --j) {
            const double xj = y[j] / D[j * 64 + j];
            y[j] = xj;
            for (int i = 0; i < j; ++i) y[i] -= xj * D[j * 64 + i];
        }
        for (int c = 0; c < 64; ++c) M.x[cb + c] = y[c];
        M.flags[b] = 1;
    }
    return true;
}

int main(int argc, char** argv)
{
    const int nf = argc > 1 ? atoi(argv[1]) : 300, span = argc > 2 ? atoi(argv[2]) : 20;
    const int max_seg = argc > 3 ? atoi(argv[3]) : 8, force = argc > 4 ? atoi(argv[4]) : 0, ctas = argc > 5 ? atoi(argv[5]) : 148;
    const int n = 6 * nf;
    unsigned long long s = 88172645463325252ull + (unsigned long long)nf * 7919u + (unsigned long long)span;
    auto rnd = [&] { s ^= s << 13, s ^= s >> 7, s ^= s << 17; return (double)(s >> 11) / 9007199254740992.0 - 0.5; };
    // block-banded SPD system in packed block form: blocks (I, J), 0 <= I - J <= span, some of them missing
    std::vector<int32_t> blocks;
    std::vector<double> Sblk;
    std::vector<double> dense((size_t)n * n, 0.0);
    for (int I = 0; I < nf; ++I)
        for (int J = std::max(0, I - span); J <= I; ++J) {
            if (I != J && I - J != span && rnd() > 0.2) continue; // keep the extreme diagonal so that span is attained
            blocks.push_back(I), blocks.push_back(J);
            for (int r = 0; r < 6; ++r)
                for (int c = 0; c < 6; ++c) {
                    double v = 0.3 * rnd();
                    if (I == J) v = r == c ? 6.0 * (span + 2) : (c < r ? v : Sblk[Sblk.size() - (size_t)(6 * r + c) + (size_t)(6 * c + r)]);
                    Sblk.push_back(v);
                    dense[(size_t)(6 * I + r) * n + 6 * J + c] = v;
                    dense[(size_t)(6 * J + c) * n + 6 * I + r] = v;
                }
        }
    std::vector<double> xt((size_t)n), bvec((size_t)n, 0.0);
    for (auto& v : xt) v = rnd();
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) bvec[(size_t)i] += dense[(size_t)i * n + j] * xt[(size_t)j];

    const SolvePlan plan = make_solve_plan(nf, blocks, ctas, max_seg, force != 0);
    const int P = plan.n_seg;
    std::vector<TaskMem> mem(plan.tasks.size());
    for (size_t t = 0; t < plan.tasks.size(); ++t) {
        const PlanTask& T = plan.tasks[t];
        if (T.n_elim % 64 || T.nrows % 64 || T.sL % 64 || T.sR % 64) return printf("FAIL: task %zu not padded to 64\n", t), 1;
        mem[t].A.assign(T.a_doubles(), NAN); // NaN outside the envelope: touching it poisons the result
        mem[t].x.assign((size_t)T.nrows, NAN);
        mem[t].flags.assign((size_t)(T.nrows / 64), 0);
        for (int r = 0; r <= T.nrows; ++r) { // k_dense_zero
            double* row = &mem[t].A[(size_t)r * T.ld];
            if (r == T.nrows) for (int c = 0; c < T.nrows; ++c) row[c] = 0.0;
            else for (int c = 64 * T.first64[(size_t)(r / 64)]; c <= r; ++c) row[c] = (r >= T.n_real && r < T.n_elim && c == r) ? 1.0 : 0.0;
        }
    }
    for (size_t k = 0; k < blocks.size() / 2; ++k) { // k_dense_assemble
        const int I = blocks[2 * k], J = blocks[2 * k + 1];
        const int tI = plan.cam_task[(size_t)I], tJ = plan.cam_task[(size_t)J], oI = plan.cam_off[(size_t)I], oJ = plan.cam_off[(size_t)J];
        int t, r0, c0;
        bool tr = false;
        if (tI == tJ) t = tI, r0 = oI, c0 = oJ;
        else if (tI == P) t = tJ, r0 = plan.tasks[(size_t)t].n_elim + plan.tasks[(size_t)t].sL + (oI - plan.sep_cols * plan.cam_sep[(size_t)I]), c0 = oJ;
        else t = tI, r0 = plan.tasks[(size_t)t].n_elim + (oJ - plan.sep_cols * plan.cam_sep[(size_t)J]), c0 = oI, tr = true;
        const PlanTask& T = plan.tasks[(size_t)t];
        for (int r = 0; r < 6; ++r)
            for (int c = 0; c < 6; ++c) {
                const double v = Sblk[36 * k + (size_t)(6 * r + c)];
                if (tr) mem[(size_t)t].A[(size_t)(r0 + c) * T.ld + c0 + r] = v;
                else mem[(size_t)t].A[(size_t)(r0 + r) * T.ld + c0 + c] = v;
            }
    }
    for (int f = 0; f < nf; ++f) {
        const PlanTask& T = plan.tasks[(size_t)plan.cam_task[(size_t)f]];
        for (int r = 0; r < 6; ++r) mem[(size_t)plan.cam_task[(size_t)f]].A[(size_t)T.nrows * T.ld + plan.cam_off[(size_t)f] + r] = bvec[(size_t)(6 * f + r)];
    }
    std::vector<std::vector<double>> Dd(plan.tasks.size());
    for (int t = 0; t < P; ++t) // phase 1
        if (!task_factor(plan.tasks[(size_t)t], mem[(size_t)t], Dd[(size_t)t])) return printf("FAIL: pivot in segment %d\n", t), 1;
    const PlanTask& top = plan.tasks[(size_t)P];
    TaskMem& tm = mem[(size_t)P];
    const int sc = plan.sep_cols;
    for (int j = 0; j + 1 < P; ++j) { // phase 2
        const PlanTask &Ta = plan.tasks[(size_t)j], &Tb = plan.tasks[(size_t)j + 1];
        const double *A = mem[(size_t)j].A.data(), *B = mem[(size_t)j + 1].A.data();
        const int ra = Ta.n_elim + Ta.sL, rb = Tb.n_elim;
        for (int r = 0; r < sc; ++r) {
            for (int c = 0; c <= r; ++c)
                tm.A[(size_t)(sc * j + r) * top.ld + sc * j + c] += A[(size_t)(ra + r) * Ta.ld + ra + c] + B[(size_t)(rb + r) * Tb.ld + rb + c];
            if (j > 0)
                for (int c = 0; c < sc; ++c) tm.A[(size_t)(sc * j + r) * top.ld + sc * (j - 1) + c] += A[(size_t)(ra + r) * Ta.ld + Ta.n_elim + c];
        }
        for (int c = 0; c < sc; ++c)
            tm.A[(size_t)top.nrows * top.ld + sc * j + c] += A[(size_t)Ta.nrows * Ta.ld + ra + c] + B[(size_t)Tb.nrows * Tb.ld + rb + c];
    }
    if (!task_factor(top, tm, Dd[(size_t)P])) return printf("FAIL: pivot in T\n"), 1; // phase 3
    if (!task_backward(top, tm, Dd[(size_t)P])) return printf("FAIL: backward of T would wait forever\n"), 1;
    for (int t = 0; t < P; ++t) { // phase 4
        const PlanTask& T = plan.tasks[(size_t)t];
        for (int i = 0; i < T.sL + T.sR; ++i) mem[(size_t)t].x[(size_t)(T.n_elim + i)] = tm.x[(size_t)(sc * (i < T.sL ? t - 1 : t) + (i < T.sL ? i : i - T.sL))];
        for (int i = 0; i < (T.sL + T.sR) / 64; ++i) mem[(size_t)t].flags[(size_t)(T.n_elim / 64 + i)] = 1;
        if (!task_backward(T, mem[(size_t)t], Dd[(size_t)t])) return printf("FAIL: backward of segment %d would wait forever\n", t), 1;
    }
    double err = 0.0;
    for (int f = 0; f < nf; ++f)
        for (int r = 0; r < 6; ++r) {
            const double x = mem[(size_t)plan.cam_task[(size_t)f]].x[(size_t)(plan.cam_off[(size_t)f] + r)];
            const double e = std::fabs(x - xt[(size_t)(6 * f + r)]);
            if (!(e <= err)) err = e; // NaN-safe max
        }
    size_t mem_d = 0;
    for (const PlanTask& T : plan.tasks) mem_d += T.a_doubles();
    printf("nf %d span %d (plan: span %d) segments %d sep_cams %d  T n %d  steps:", nf, span, plan.span, P, plan.sep_cams, top.n_elim);
    for (const PlanTask& T : plan.tasks) printf(" %d", T.n_elim / 64);
    printf("  matrices %.1f MB  max |x - x_true| %.3e\n", mem_d * 8e-6, err);
    return (err < 1e-9) ? 0 : 1;
}
