// synth.cpp — deterministic synthetic bundle-adjustment problems.
//
// Host-only, built as its own library (libviso_synth.so: no CUDA, so the CPU arm of bench.py never maps the
// CUDA backend).  One generator feeds both the CUDA backend and the CPU oracle so that
// they read identical bytes.  Scene conventions follow the reference:
//   intrinsics fx=517.3 fy=516.5 cx=325.1 cy=249.7, 640x480   src/main.cpp:14-17, :73
//   mean scene depth 1                                         src/viso.cpp:622-637
//   4 px image margin (half_patch_size)                        include/viso.h:25
//   measurements are float32 (cv::Point2f)                     include/keyframe.h:15
//   frame 0 is the identity pose                               include/keyframe.h:34-35
#include "../../include/viso_ba.h" // status codes only
#include "../../include/viso_ba_synth.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <unordered_map>
#include <vector>

namespace {

struct Rng { // splitmix64-seeded xoshiro256**
    uint64_t s[4];
    explicit Rng(uint64_t seed)
    {
        uint64_t x = seed;
        for (int i = 0; i < 4; ++i) {
            x += 0x9E3779B97F4A7C15ull;
            uint64_t z = x;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            s[i] = z ^ (z >> 31);
        }
    }
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t next()
    {
        const uint64_t r = rotl(s[1] * 5, 7) * 9;
        const uint64_t t = s[1] << 17;
        s[2] ^= s[0];
        s[3] ^= s[1];
        s[1] ^= s[2];
        s[0] ^= s[3];
        s[2] ^= t;
        s[3] = rotl(s[3], 45);
        return r;
    }
    double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
    double normal()
    { // Box-Muller, one value per call (the sine branch is discarded for simplicity)
        double u1 = uniform();
        if (u1 < 1e-300) u1 = 1e-300;
        const double u2 = uniform();
        return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586476925 * u2);
    }
};

struct Pose {
    double q[4]; // x y z w
    double t[3];
};

void quat_to_R(const double q[4], double R[9])
{
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    R[0] = 1 - 2 * (y * y + z * z);
    R[1] = 2 * (x * y - z * w);
    R[2] = 2 * (x * z + y * w);
    R[3] = 2 * (x * y + z * w);
    R[4] = 1 - 2 * (x * x + z * z);
    R[5] = 2 * (y * z - x * w);
    R[6] = 2 * (x * z - y * w);
    R[7] = 2 * (y * z + x * w);
    R[8] = 1 - 2 * (x * x + y * y);
}

// left perturbation pose <- exp((v, w)) * pose, closed form (no small-angle branch needed
// below ~1e-8 because sigma is ~1e-2; a series guards the exact-zero case)
void perturb(Pose& p, const double v[3], const double w[3])
{
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    const double th = std::sqrt(th2);
    double im, re, a, b;
    if (th2 < 1e-10) {
        im = 0.5 - th2 / 48.0;
        re = 1.0 - th2 / 8.0;
        a = 0.5 - th2 / 24.0;
        b = 1.0 / 6.0 - th2 / 120.0;
    } else {
        im = std::sin(0.5 * th) / th;
        re = std::cos(0.5 * th);
        a = (1.0 - std::cos(th)) / th2;
        b = (th - std::sin(th)) / (th2 * th);
    }
    const double dq[4] = { im * w[0], im * w[1], im * w[2], re };
    const double c1[3] = { w[1] * v[2] - w[2] * v[1], w[2] * v[0] - w[0] * v[2],
        w[0] * v[1] - w[1] * v[0] };
    const double c2[3] = { w[1] * c1[2] - w[2] * c1[1], w[2] * c1[0] - w[0] * c1[2],
        w[0] * c1[1] - w[1] * c1[0] };
    const double dt[3] = { v[0] + a * c1[0] + b * c2[0], v[1] + a * c1[1] + b * c2[1],
        v[2] + a * c1[2] + b * c2[2] };
    double dR[9];
    quat_to_R(dq, dR);
    const double bx = p.q[0], by = p.q[1], bz = p.q[2], bw = p.q[3];
    double q[4] = { dq[3] * bx + dq[0] * bw + dq[1] * bz - dq[2] * by,
        dq[3] * by - dq[0] * bz + dq[1] * bw + dq[2] * bx,
        dq[3] * bz + dq[0] * by - dq[1] * bx + dq[2] * bw,
        dq[3] * bw - dq[0] * bx - dq[1] * by - dq[2] * bz };
    const double n = 1.0 / std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int i = 0; i < 4; ++i) p.q[i] = q[i] * n;
    const double t0 = p.t[0], t1 = p.t[1], t2 = p.t[2];
    p.t[0] = dR[0] * t0 + dR[1] * t1 + dR[2] * t2 + dt[0];
    p.t[1] = dR[3] * t0 + dR[4] * t1 + dR[5] * t2 + dt[1];
    p.t[2] = dR[6] * t0 + dR[7] * t1 + dR[8] * t2 + dt[2];
}

} // namespace

struct viso_ba_synth {
    viso_ba_synth_spec spec;
    std::vector<double> poses_true, poses_init, points_true, points_init;
    std::vector<int32_t> edge_pose, edge_point;
    std::vector<float> edge_uv;
};

extern "C" int32_t viso_ba_synth_preset(int32_t config_id, int32_t window, viso_ba_synth_spec* o)
{
    if (!o) return VISO_BA_ERR_ARG;
    std::memset(o, 0, sizeof(*o));
    o->struct_size = (int32_t)sizeof(*o);
    o->fx = 517.3; // src/main.cpp:14-17
    o->fy = 516.5;
    o->cx = 325.1;
    o->cy = 249.7;
    o->width = 640.0; // src/main.cpp:73
    o->height = 480.0;
    o->margin = 4.0; // include/viso.h:25
    o->baseline = 0.05;
    o->noise_px = 1.0;
    o->outlier_frac = 0.0;
    o->pose_rot_sigma = 0.01;
    o->pose_trans_sigma = 0.02;
    o->point_sigma = 0.02;
    o->shuffle_edges = 1;
    switch (config_id) {
    case 1:
        o->layout = 0, o->n_poses = 10, o->n_points = 2000, o->target_edges = 15000;
        o->seed = 0x5EED0001ull;
        break;
    case 2:
        o->layout = 0, o->n_poses = 20, o->n_points = 10000, o->target_edges = 80000;
        o->outlier_frac = 0.05;
        o->seed = 0x5EED0002ull;
        break;
    case 3:
        o->layout = 1, o->n_poses = 200, o->n_points = 100000, o->target_edges = 1000000;
        o->baseline = 0.25;
        o->seed = 0x5EED0003ull;
        break;
    case 4:
        o->layout = 2, o->n_poses = 2000, o->n_points = 1000000, o->target_edges = 10000000;
        o->baseline = 0.1;
        o->seed = 0x5EED0004ull;
        break;
    case 5:
        o->layout = 0, o->n_poses = 10, o->n_points = 2000, o->target_edges = 15000;
        o->seed = 0x5EED0500ull + (uint64_t)(window < 0 ? 0 : window);
        break;
    default:
        return VISO_BA_ERR_ARG;
    }
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_synth_create(const viso_ba_synth_spec* sp, viso_ba_synth** out)
{
    if (!sp || !out || sp->struct_size != (int32_t)sizeof(viso_ba_synth_spec)) return VISO_BA_ERR_ARG;
    if (sp->n_poses < 1 || sp->n_points < 1 || sp->target_edges < 1) return VISO_BA_ERR_ARG;
    if (sp->n_poses > 0x7fffffff || sp->n_points > 0x7fffffff) return VISO_BA_ERR_ARG;
    viso_ba_synth* s = new (std::nothrow) viso_ba_synth();
    if (!s) return VISO_BA_ERR_CAPACITY;
    s->spec = *sp;
    const int64_t Nc = sp->n_poses, Np = sp->n_points;
    Rng rng(sp->seed);

    // ---- true camera poses (world -> camera) -------------------------------------------
    std::vector<Pose> cams((size_t)Nc);
    std::vector<double> centre((size_t)Nc * 3), Rwc((size_t)Nc * 9);
    int64_t grid_w = 1;
    if (sp->layout == 1) {
        grid_w = (int64_t)std::ceil(std::sqrt(2.0 * (double)Nc));
        if (grid_w < 1) grid_w = 1;
    }
    for (int64_t i = 0; i < Nc; ++i) {
        double C[3];
        const double di = (double)i;
        if (sp->layout == 1) {
            const int64_t gx = i % grid_w, gy = i / grid_w;
            C[0] = sp->baseline * (double)gx + 0.02 * std::sin(di);
            C[1] = sp->baseline * (double)gy + 0.02 * (std::cos(di) - 1.0);
            C[2] = 0.01 * std::sin(0.7 * di);
        } else {
            C[0] = sp->baseline * di;
            C[1] = 0.02 * std::sin(di);
            C[2] = 0.02 * (std::cos(di) - 1.0);
        }
        // small yaw about the camera y axis, bounded so long trajectories keep looking forward
        const double yaw = (sp->layout == 0) ? 0.01 * di : 0.05 * std::sin(0.37 * di);
        Pose p;
        p.q[0] = 0.0;
        p.q[1] = std::sin(0.5 * yaw);
        p.q[2] = 0.0;
        p.q[3] = std::cos(0.5 * yaw);
        double R[9];
        quat_to_R(p.q, R);
        for (int r = 0; r < 3; ++r) p.t[r] = -(R[3 * r] * C[0] + R[3 * r + 1] * C[1] + R[3 * r + 2] * C[2]);
        cams[(size_t)i] = p;
        for (int r = 0; r < 3; ++r) centre[(size_t)i * 3 + r] = C[r];
        for (int r = 0; r < 9; ++r) Rwc[(size_t)i * 9 + r] = R[r];
    }

    // ---- spatial hash of camera centres (cell = 1 scene unit) ---------------------------
    std::unordered_map<int64_t, std::vector<int32_t>> cells;
    auto cell_key = [](int64_t ix, int64_t iy) { return ix * 0x100000 + iy; };
    for (int64_t i = 0; i < Nc; ++i) {
        const int64_t ix = (int64_t)std::floor(centre[(size_t)i * 3]);
        const int64_t iy = (int64_t)std::floor(centre[(size_t)i * 3 + 1]);
        cells[cell_key(ix, iy)].push_back((int32_t)i);
    }

    // ---- points: back-project a random pixel of a random camera at depth in [0.5, 1.5] ---
    s->points_true.resize((size_t)Np * 3);
    std::vector<int64_t> vis_ptr((size_t)Np + 1, 0);
    std::vector<int32_t> vis_cam;
    vis_cam.reserve((size_t)Np * 12);
    std::vector<int32_t> cand;
    const double umin = sp->margin, umax = sp->width - sp->margin;
    const double vmin = sp->margin, vmax = sp->height - sp->margin;
    for (int64_t p = 0; p < Np; ++p) {
        for (int attempt = 0;; ++attempt) {
            const int64_t c = (int64_t)(rng.uniform() * (double)Nc) % Nc;
            const double u = umin + (umax - umin) * rng.uniform();
            const double v = vmin + (vmax - vmin) * rng.uniform();
            const double z = 0.5 + rng.uniform();
            const double Pc[3] = { (u - sp->cx) / sp->fx * z, (v - sp->cy) / sp->fy * z, z };
            const double* R = &Rwc[(size_t)c * 9];
            const double* C = &centre[(size_t)c * 3];
            double Pw[3]; // Pw = R^T (Pc - t) = R^T Pc + C
            for (int r = 0; r < 3; ++r) Pw[r] = R[r] * Pc[0] + R[3 + r] * Pc[1] + R[6 + r] * Pc[2] + C[r];
            cand.clear();
            const int64_t ix = (int64_t)std::floor(Pw[0]), iy = (int64_t)std::floor(Pw[1]);
            for (int64_t dx = -2; dx <= 2; ++dx)
                for (int64_t dy = -2; dy <= 2; ++dy) {
                    auto it = cells.find(cell_key(ix + dx, iy + dy));
                    if (it == cells.end()) continue;
                    for (int32_t k : it->second) {
                        const double* Rk = &Rwc[(size_t)k * 9];
                        const double* tk = cams[(size_t)k].t;
                        const double x = Rk[0] * Pw[0] + Rk[1] * Pw[1] + Rk[2] * Pw[2] + tk[0];
                        const double y = Rk[3] * Pw[0] + Rk[4] * Pw[1] + Rk[5] * Pw[2] + tk[1];
                        const double zz = Rk[6] * Pw[0] + Rk[7] * Pw[1] + Rk[8] * Pw[2] + tk[2];
                        if (zz < 0.2) continue;
                        const double uu = sp->fx * x / zz + sp->cx, vv = sp->fy * y / zz + sp->cy;
                        if (uu >= umin && uu < umax && vv >= vmin && vv < vmax) cand.push_back(k);
                    }
                }
            if (cand.size() >= 2 || (Nc < 2 && !cand.empty()) || attempt > 64) {
                std::sort(cand.begin(), cand.end());
                for (int r = 0; r < 3; ++r) s->points_true[(size_t)p * 3 + r] = Pw[r];
                vis_cam.insert(vis_cam.end(), cand.begin(), cand.end());
                vis_ptr[(size_t)p + 1] = (int64_t)vis_cam.size();
                break;
            }
        }
    }

    // ---- Bernoulli thinning to the target edge count, every point keeps >= 2 -----------
    const double keep = std::min(1.0, (double)sp->target_edges / (double)std::max<size_t>(1, vis_cam.size()));
    std::vector<int32_t> e_pose, e_point;
    e_pose.reserve((size_t)sp->target_edges + (size_t)Np);
    e_point.reserve((size_t)sp->target_edges + (size_t)Np);
    std::vector<int32_t> kept;
    for (int64_t p = 0; p < Np; ++p) {
        kept.clear();
        const int64_t a = vis_ptr[(size_t)p], b = vis_ptr[(size_t)p + 1];
        for (int64_t k = a; k < b; ++k)
            if (rng.uniform() < keep) kept.push_back(vis_cam[(size_t)k]);
        for (int64_t k = a; k < b && kept.size() < 2; ++k) { // force the first ones back in
            const int32_t c = vis_cam[(size_t)k];
            if (std::find(kept.begin(), kept.end(), c) == kept.end()) kept.push_back(c);
        }
        for (int32_t c : kept) {
            e_pose.push_back(c);
            e_point.push_back((int32_t)p);
        }
    }
    const int64_t Ne = (int64_t)e_pose.size();

    // ---- measurements: projection + N(0, noise^2), outliers 20..50 px away, float32 ----
    std::vector<float> uv((size_t)Ne * 2);
    for (int64_t e = 0; e < Ne; ++e) {
        const int32_t k = e_pose[(size_t)e];
        const double* Pw = &s->points_true[(size_t)e_point[(size_t)e] * 3];
        const double* Rk = &Rwc[(size_t)k * 9];
        const double* tk = cams[(size_t)k].t;
        const double x = Rk[0] * Pw[0] + Rk[1] * Pw[1] + Rk[2] * Pw[2] + tk[0];
        const double y = Rk[3] * Pw[0] + Rk[4] * Pw[1] + Rk[5] * Pw[2] + tk[1];
        const double z = Rk[6] * Pw[0] + Rk[7] * Pw[1] + Rk[8] * Pw[2] + tk[2];
        double u = sp->fx * x / z + sp->cx + sp->noise_px * rng.normal();
        double v = sp->fy * y / z + sp->cy + sp->noise_px * rng.normal();
        if (sp->outlier_frac > 0.0 && rng.uniform() < sp->outlier_frac) {
            const double ang = 6.283185307179586476925 * rng.uniform();
            const double rad = 20.0 + 30.0 * rng.uniform();
            u += rad * std::cos(ang);
            v += rad * std::sin(ang);
        }
        uv[(size_t)e * 2] = (float)u;
        uv[(size_t)e * 2 + 1] = (float)v;
    }

    // ---- initial state: perturbed truth, pose 0 untouched -------------------------------
    s->poses_true.resize((size_t)Nc * 7);
    s->poses_init.resize((size_t)Nc * 7);
    for (int64_t i = 0; i < Nc; ++i) {
        Pose p = cams[(size_t)i];
        for (int r = 0; r < 4; ++r) s->poses_true[(size_t)i * 7 + r] = p.q[r];
        for (int r = 0; r < 3; ++r) s->poses_true[(size_t)i * 7 + 4 + r] = p.t[r];
        if (i > 0) {
            double v[3], w[3];
            for (int r = 0; r < 3; ++r) w[r] = sp->pose_rot_sigma * rng.normal();
            for (int r = 0; r < 3; ++r) v[r] = sp->pose_trans_sigma * rng.normal();
            perturb(p, v, w);
        }
        for (int r = 0; r < 4; ++r) s->poses_init[(size_t)i * 7 + r] = p.q[r];
        for (int r = 0; r < 3; ++r) s->poses_init[(size_t)i * 7 + 4 + r] = p.t[r];
    }
    s->points_init.resize((size_t)Np * 3);
    for (size_t i = 0; i < (size_t)Np * 3; ++i)
        s->points_init[i] = s->points_true[i] + sp->point_sigma * rng.normal();

    // ---- optional deterministic shuffle of the edge order -------------------------------
    std::vector<int64_t> order((size_t)Ne);
    for (int64_t e = 0; e < Ne; ++e) order[(size_t)e] = e;
    if (sp->shuffle_edges) {
        for (int64_t e = Ne - 1; e > 0; --e) {
            const int64_t j = (int64_t)(rng.next() % (uint64_t)(e + 1));
            std::swap(order[(size_t)e], order[(size_t)j]);
        }
    }
    s->edge_pose.resize((size_t)Ne);
    s->edge_point.resize((size_t)Ne);
    s->edge_uv.resize((size_t)Ne * 2);
    for (int64_t e = 0; e < Ne; ++e) {
        const int64_t src = order[(size_t)e];
        s->edge_pose[(size_t)e] = e_pose[(size_t)src];
        s->edge_point[(size_t)e] = e_point[(size_t)src];
        s->edge_uv[(size_t)e * 2] = uv[(size_t)src * 2];
        s->edge_uv[(size_t)e * 2 + 1] = uv[(size_t)src * 2 + 1];
    }
    *out = s;
    return VISO_BA_OK;
}

extern "C" void viso_ba_synth_destroy(viso_ba_synth* s) { delete s; }

extern "C" int64_t viso_ba_synth_num_edges(const viso_ba_synth* s)
{
    return s ? (int64_t)s->edge_pose.size() : -1;
}

extern "C" int32_t viso_ba_synth_fetch(const viso_ba_synth* s, double* poses_init7,
    double* poses_true7, double* points_init3, double* points_true3, int32_t* edge_pose,
    int32_t* edge_point, float* edge_uv)
{
    if (!s) return VISO_BA_ERR_ARG;
    if (poses_init7) std::memcpy(poses_init7, s->poses_init.data(), s->poses_init.size() * sizeof(double));
    if (poses_true7) std::memcpy(poses_true7, s->poses_true.data(), s->poses_true.size() * sizeof(double));
    if (points_init3) std::memcpy(points_init3, s->points_init.data(), s->points_init.size() * sizeof(double));
    if (points_true3) std::memcpy(points_true3, s->points_true.data(), s->points_true.size() * sizeof(double));
    if (edge_pose) std::memcpy(edge_pose, s->edge_pose.data(), s->edge_pose.size() * sizeof(int32_t));
    if (edge_point) std::memcpy(edge_point, s->edge_point.data(), s->edge_point.size() * sizeof(int32_t));
    if (edge_uv) std::memcpy(edge_uv, s->edge_uv.data(), s->edge_uv.size() * sizeof(float));
    return VISO_BA_OK;
}
