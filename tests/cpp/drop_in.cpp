// drop_in.cpp — a caller written against the reference's class surface
// (reference include/bundle_adjuster.h:111, :169-173) built against this repo's header.
// usage: drop_in [--optimize]   (without the flag: staging only, needs no GPU)
#include "bundle_adjuster.h"

#include <cmath>
#include <cstdio>
#include <cstring>

static void project(const double fx, const double fy, const double cx, const double cy, const double* t, const double* P, double* uv)
{ // identity rotation; keyframe.h:82-89
    const double x = P[0] + t[0], y = P[1] + t[1], z = P[2] + t[2];
    uv[0] = fx * x / z + cx;
    uv[1] = fy * y / z + cy;
}

int main(int argc, char** argv)
{
    const bool optimize = argc > 1 && !std::strcmp(argv[1], "--optimize");
    BundleAdjuster ba; // default pools of 1000, as the reference
    const double fx = 517.3, fy = 516.5, cx = 325.1, cy = 249.7;
    ba.SetIntrinsics(fx, fy, cx, cy);
    const int ncam = 4, npt = 60;
    double cam_t[ncam][3];
    for (int c = 0; c < ncam; ++c) {
        const double R[9] = { 1, 0, 0, 0, 1, 0, 0, 0, 1 };
        cam_t[c][0] = -0.05 * c, cam_t[c][1] = 0.01 * c, cam_t[c][2] = 0.0;
        // perturbed start for every pose but the first
        const double t0[3] = { cam_t[c][0] + (c ? 0.004 : 0.0), cam_t[c][1] - (c ? 0.003 : 0.0), cam_t[c][2] + (c ? 0.002 : 0.0) };
        if (ba.AddPose(R, t0) != c) return 10;
    }
    int edges = 0;
    for (int p = 0; p < npt; ++p) {
        const double P[3] = { -0.3 + 0.6 * ((p * 37) % 60) / 60.0, -0.2 + 0.4 * ((p * 11) % 60) / 60.0, 0.8 + 0.5 * ((p * 23) % 60) / 60.0 };
        const double P0[3] = { P[0] + 0.01 * std::sin(p), P[1] - 0.01 * std::cos(p), P[2] + 0.01 * std::sin(2.0 * p) };
        if (ba.AddPoint(P0) != p) return 11;
        for (int c = 0; c < ncam; ++c) {
            double uv[2];
            project(fx, fy, cx, cy, cam_t[c], P, uv);
            ba.AddEdge(p, c, uv[0], uv[1]);
            if (ba.LastStatus() != VISO_BA_OK) return 12;
            ++edges;
        }
    }
    if (ba.NumPoints() != npt || ba.NumPoses() != ncam || ba.NumEdges() != edges) return 13;
    // error behaviour: bad index is reported, never exit()s (reference :162-165 exits)
    ba.AddEdge(npt + 5, 0, 1.0, 1.0);
    if (ba.LastStatus() != VISO_BA_ERR_INDEX) return 14;
    // the reference's two-argument AddEdge still compiles; without a pixel Optimize refuses
    ba.AddEdge(0, 0);
    if (ba.Optimize(1) != VISO_BA_ERR_STATE) return 15;
    double uv[2];
    const double P0true[3] = { -0.3, -0.2, 0.8 };
    project(fx, fy, cx, cy, cam_t[0], P0true, uv);
    ba.SetEdgeMeasurement(0, uv[0], uv[1]); // a duplicate (point 0, pose 0) edge: allowed
    if (ba.LastStatus() != VISO_BA_OK) return 16;

    viso_ba_summary s;
    const int st = ba.Optimize(15, &s);
    if (!optimize) { // no GPU: must fail loudly with ERR_CUDA, never compute on the CPU
        std::printf("staging ok, optimize status %d (%s)\n", st, ba.LastError());
        return st == VISO_BA_ERR_CUDA || st == VISO_BA_OK ? 0 : 17;
    }
    if (st != VISO_BA_OK) {
        std::printf("optimize failed: %s\n", ba.LastError());
        return 18;
    }
    double worst = 0.0;
    for (int c = 0; c < ncam; ++c) {
        double q[7];
        ba.GetPose(c, q);
        for (int k = 0; k < 3; ++k) worst = std::fmax(worst, std::fabs(q[4 + k] - cam_t[c][k]));
    }
    std::printf("cost %.3e -> %.3e in %d iterations, worst translation error %.3e\n", s.initial_cost, s.final_cost, s.iterations, worst);
    // noise-free measurements (rounded to float32): the truth is recovered up to that rounding
    return (s.final_cost < 1e-4 * s.initial_cost && worst < 1e-4) ? 0 : 19;
}
