// ba_solve_plan.h — host-side plan of the large reduced-system solve (ba_dense.cu): how the free poses of ONE problem
// are cut into interior segments and separators, where every stored S block lands, and which rows each elimination
// step touches.  Pure C++ (no CUDA) so that the plan can be tested on the CPU (tests/cpp/solve_plan_test.cpp runs the
// same plan through a sequential emulation of the kernel and compares with a plain dense solve).
//
// Reference concept: g2o LinearSolverDense on the Schur-reduced pose matrix (reference include/bundle_adjuster.h:113-114).
// The reference factorises one dense matrix on one core; this plan is what lets 148 SMs (and, through the replicated
// solve, every rank of a sharded problem) share a banded one: see the comment on CholTask in ba_device.h.
#ifndef VISO_BA_SOLVE_PLAN_H
#define VISO_BA_SOLVE_PLAN_H

#include <algorithm>
#include <cstdint>
#include <vector>

namespace visoba {

struct PlanTask { // mirrors CholTask without the pointers
    int32_t n_elim = 0, nrows = 0, n_real = 0, ld = 0, sL = 0, sR = 0, cta0 = 0, ncta = 0;
    std::vector<int32_t> first64;  // [nrows / 64]
    std::vector<int32_t> steps;    // 4 per elimination step: r1_end, r2_begin, r2_end, 0
    size_t a_doubles() const { return (size_t)(nrows + 1) * (size_t)ld; }
    size_t dd_doubles() const { return (size_t)(n_elim / 64) * (64 * 64 + 128); } // kCholDdStride (ba_device.h) per block
};

struct SolvePlan {
    int32_t nf = 0, n_seg = 0, sep_cams = 0, sep_cols = 0, span = 0;
    std::vector<PlanTask> tasks;     // n_seg interior tasks, then T
    std::vector<int32_t> cam_task, cam_off, cam_sep; // per free pose
    std::vector<int32_t> sep_cam0;   // first free pose of separator j
    std::vector<int32_t> row_ptr;    // prefix of nrows + 1
    int32_t total_rows = 0;
};

inline int32_t round64(int32_t x) { return (x + 63) & ~63; }

// steps of a task whose rows [0, n_band) carry a monotone envelope `first64` (first column block of every row block):
// step k touches the row blocks (k, last64[k]] of the band, last64[k] = last row block whose envelope reaches column block k
inline void band_steps(const std::vector<int32_t>& first64, int32_t n_band_blocks, int32_t n_steps, std::vector<int32_t>& r1_end)
{
    r1_end.assign((size_t)n_steps, 0);
    for (int32_t k = 0, i = 0; k < n_steps; ++k) {
        if (i < k) i = k;
        while (i + 1 < n_band_blocks && first64[(size_t)i + 1] <= k) ++i;
        r1_end[(size_t)k] = 64 * (std::max(i, k) + 1);
    }
}

// blocks: (I, J) pairs of free-pose indices, I >= J, of every stored S block (diagonal blocks included).
// max_seg: upper bound on the number of interior segments (<= 1: no partition); force: take max_seg segments whenever the
// structure allows it (tests), otherwise only when every interior is at least twice as wide as a separator.
inline SolvePlan make_solve_plan(int32_t nf, const std::vector<int32_t>& blocks, int32_t n_cta, int32_t max_seg, bool force)
{
    SolvePlan p;
    p.nf = nf;
    const size_t nb = blocks.size() / 2;
    int32_t span = 0;
    for (size_t k = 0; k < nb; ++k) span = std::max(span, blocks[2 * k] - blocks[2 * k + 1]);
    p.span = span;
    const int32_t sep_cams = 32 * std::max(1, (span + 31) / 32); // 32 poses = 192 columns = 3 blocks of 64
    int32_t P = 0;
    if (max_seg >= 2) {
        // interior i must hold >= 1 pose; by default it should also dwarf the separators or the partition does not pay
        const int32_t min_int = force ? 1 : 2 * sep_cams;
        for (int32_t q = std::min(max_seg, std::max(1, n_cta)); q >= 2; --q)
            if ((int64_t)nf - (int64_t)(q - 1) * sep_cams >= (int64_t)q * min_int) {
                P = q;
                break;
            }
    }
    p.cam_task.assign((size_t)nf, 0), p.cam_off.assign((size_t)nf, 0), p.cam_sep.assign((size_t)nf, -1);
    auto finish_rows = [&] {
        p.row_ptr.assign(p.tasks.size() + 1, 0);
        for (size_t t = 0; t < p.tasks.size(); ++t) p.row_ptr[t + 1] = p.row_ptr[t] + p.tasks[t].nrows + 1;
        p.total_rows = p.row_ptr.back();
    };
    auto unpartitioned = [&] {
        p.n_seg = 0, p.sep_cams = 0, p.sep_cols = 0;
        p.tasks.assign(1, PlanTask());
        PlanTask& T = p.tasks[0];
        T.n_real = 6 * nf, T.n_elim = T.nrows = round64(6 * nf), T.ld = T.nrows + 8, T.cta0 = 0, T.ncta = n_cta;
        const int32_t nblk = T.nrows / 64;
        T.first64.resize((size_t)nblk);
        for (int32_t i = 0; i < nblk; ++i) T.first64[(size_t)i] = i;
        for (size_t k = 0; k < nb; ++k) {
            const int32_t I = blocks[2 * k], J = blocks[2 * k + 1], cb = 6 * J / 64;
            for (int32_t rb = 6 * I / 64; rb <= (6 * I + 5) / 64; ++rb) T.first64[(size_t)rb] = std::min(T.first64[(size_t)rb], cb);
        }
        for (int32_t i = nblk - 2; i >= 0; --i) T.first64[(size_t)i] = std::min(T.first64[(size_t)i], T.first64[(size_t)i + 1]);
        std::vector<int32_t> r1;
        band_steps(T.first64, nblk, nblk, r1);
        T.steps.assign(4 * (size_t)nblk, 0);
        for (int32_t k = 0; k < nblk; ++k) T.steps[4 * (size_t)k] = r1[(size_t)k];
        for (int32_t f = 0; f < nf; ++f) p.cam_task[(size_t)f] = 0, p.cam_off[(size_t)f] = 6 * f, p.cam_sep[(size_t)f] = -1;
        finish_rows();
    };
    if (P < 2) {
        unpartitioned();
        return p;
    }
    // ---- segments and separators along the pose chain: seg 0, sep 0, seg 1, sep 1, ..., seg P-1 ----
    p.n_seg = P, p.sep_cams = sep_cams, p.sep_cols = 6 * sep_cams;
    const int32_t interior_total = nf - (P - 1) * sep_cams, base = interior_total / P, extra = interior_total % P;
    std::vector<int32_t> seg0((size_t)P + 1, 0), segn((size_t)P, 0);
    p.sep_cam0.assign((size_t)std::max(P - 1, 0), 0);
    {
        int32_t f = 0;
        for (int32_t i = 0; i < P; ++i) {
            seg0[(size_t)i] = f, segn[(size_t)i] = base + (i < extra ? 1 : 0);
            for (int32_t c = 0; c < segn[(size_t)i]; ++c, ++f) p.cam_task[(size_t)f] = i, p.cam_off[(size_t)f] = 6 * c;
            if (i + 1 < P) {
                p.sep_cam0[(size_t)i] = f;
                for (int32_t c = 0; c < sep_cams; ++c, ++f)
                    p.cam_task[(size_t)f] = P, p.cam_off[(size_t)f] = p.sep_cols * i + 6 * c, p.cam_sep[(size_t)f] = i;
            }
        }
    }
    p.tasks.assign((size_t)P + 1, PlanTask());
    const int32_t per = std::max(1, n_cta / P);
    for (int32_t i = 0; i < P; ++i) {
        PlanTask& t = p.tasks[(size_t)i];
        t.n_real = 6 * segn[(size_t)i], t.n_elim = round64(t.n_real);
        t.sL = i > 0 ? p.sep_cols : 0, t.sR = i + 1 < P ? p.sep_cols : 0;
        t.nrows = t.n_elim + t.sL + t.sR, t.ld = t.nrows + 8;
        t.cta0 = std::min(i * per, std::max(0, n_cta - 1)), t.ncta = std::min(per, n_cta - t.cta0);
        t.first64.resize((size_t)(t.nrows / 64));
        for (int32_t r = 0; r < t.nrows / 64; ++r) t.first64[(size_t)r] = r < t.n_elim / 64 ? r : 0;
    }
    PlanTask& T = p.tasks[(size_t)P];
    T.n_real = T.n_elim = T.nrows = p.sep_cols * (P - 1), T.ld = T.nrows + 8, T.cta0 = 0, T.ncta = n_cta;
    // ---- where every stored block lands; envelope of the interiors; first coupled column of the right separators ----
    std::vector<int32_t> fR((size_t)P, 1 << 28);
    bool ok = true;
    for (size_t k = 0; k < nb && ok; ++k) {
        const int32_t I = blocks[2 * k], J = blocks[2 * k + 1];
        const int32_t tI = p.cam_task[(size_t)I], tJ = p.cam_task[(size_t)J];
        if (tI < P && tJ < P) { // interior x interior: the same segment, or the separator is too narrow
            if (tI != tJ) ok = false;
            else {
                PlanTask& t = p.tasks[(size_t)tI];
                const int32_t cb = p.cam_off[(size_t)J] / 64;
                for (int32_t rb = p.cam_off[(size_t)I] / 64; rb <= (p.cam_off[(size_t)I] + 5) / 64; ++rb)
                    t.first64[(size_t)rb] = std::min(t.first64[(size_t)rb], cb);
            }
        } else if (tI == P && tJ == P) { // separator x separator: the same one or neighbours (T is block tridiagonal)
            if (p.cam_sep[(size_t)I] - p.cam_sep[(size_t)J] > 1) ok = false;
        } else if (tI == P) { // I in a separator, J interior and BEFORE it in the chain: J's segment has it on its right
            if (p.cam_sep[(size_t)I] != tJ) ok = false;
            else fR[(size_t)tJ] = std::min(fR[(size_t)tJ], p.cam_off[(size_t)J] / 64);
        } else { // I interior, J in a separator before it: I's segment has that separator on its left
            if (p.cam_sep[(size_t)J] != tI - 1) ok = false;
        }
    }
    if (!ok) { // cannot happen while sep_cams >= span; kept as a guard
        unpartitioned();
        return p;
    }
    for (int32_t i = 0; i < P; ++i) {
        PlanTask& t = p.tasks[(size_t)i];
        const int32_t nbi = t.n_elim / 64;
        for (int32_t r = nbi - 2; r >= 0; --r) t.first64[(size_t)r] = std::min(t.first64[(size_t)r], t.first64[(size_t)r + 1]);
        const int32_t fr = std::min(fR[(size_t)i], nbi); // nbi: the right separator never couples (kept out of every step)
        for (int32_t r = 0; r < t.sR / 64; ++r) t.first64[(size_t)(nbi + t.sL / 64 + r)] = std::min(fr, nbi);
        std::vector<int32_t> r1;
        band_steps(t.first64, nbi, nbi, r1);
        t.steps.assign(4 * (size_t)nbi, 0);
        for (int32_t k = 0; k < nbi; ++k) {
            const bool right = t.sR > 0 && k >= fr;
            t.steps[4 * (size_t)k] = r1[(size_t)k];
            if (t.sL > 0 || right) {
                t.steps[4 * (size_t)k + 1] = t.sL > 0 ? t.n_elim : t.n_elim + t.sL;
                t.steps[4 * (size_t)k + 2] = t.n_elim + t.sL + (right ? t.sR : 0);
            }
        }
    }
    { // T: block tridiagonal in separators
        const int32_t nblk = T.nrows / 64, bps = p.sep_cols / 64;
        T.first64.resize((size_t)nblk);
        for (int32_t r = 0; r < nblk; ++r) T.first64[(size_t)r] = bps * std::max(0, r / bps - 1);
        std::vector<int32_t> r1;
        band_steps(T.first64, nblk, nblk, r1);
        T.steps.assign(4 * (size_t)nblk, 0);
        for (int32_t k = 0; k < nblk; ++k) T.steps[4 * (size_t)k] = r1[(size_t)k];
    }
    finish_rows();
    return p;
}

} // namespace visoba
#endif
