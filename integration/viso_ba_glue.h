// viso_ba_glue.h — the C++ wiring of the B200 bundle adjuster into viso's VO loop.
//
// What the reference would need at its two call sites and does not have (nothing there calls the adjuster):
//   * after two-view initialisation, reference src/viso.cpp:77-98 (2 keyframes + triangulated points + observations),
//   * in the tracking branch after LKAlignment, reference src/viso.cpp:113-138 (sliding window of the last N keyframes),
// namely: pull (point, pose, pixel) triples out of Map / MapPoint / Keyframe (reference include/map.h:15-19,
// include/map_point.h:21-31, include/keyframe.h:100-106), keep the last N keyframes in a RingBuffer-shaped window
// (reference include/ring_buffer.h:10-42 is included by viso.h:8 and never instantiated), call Optimize, and write the
// refined poses and points back (Keyframe::SetR / SetT exist, keyframe.h:105-106; MapPoint has NO setter for Pw_,
// map_point.h:31 — the one-line `SetWorldPos` below is the change a maintainer adds).
//
// Everything is a template over the reference's own accessor NAMES (Keyframes(), GetPoints(), GetObservations(),
// GetWorldPos(), Keypoints()[i].pt.x, GetR()(r, c), GetT()(i), SetR, SetT), so the same text compiles
//   * inside viso, against the real Keyframe / MapPoint / Map with Eigen, Sophus and OpenCV (`ViSoTypes` below is
//     enabled by __has_include(<opencv2/core.hpp>) and <Eigen/Core>), and
//   * here, where none of those libraries exist, against the plain structs of tests/cpp/glue_test.cpp, which is how
//     tests/test_cpp_glue.py compiles and RUNS it (extraction on the CPU, the whole chain on the GPU).
#ifndef VISO_BA_GLUE_H
#define VISO_BA_GLUE_H

#include "../include/bundle_adjuster.h"

#include <cmath>
#include <cstdint>
#include <map>
#include <vector>

namespace viso_ba_glue {

// Fixed-capacity FIFO of the last N items, oldest first — the interface of the reference's RingBuffer<N, T>
// (Push / GetSize / operator[] by value, ring_buffer.h:19-36) so that either class can hold the window.
template <int N, typename T> class Window {
public:
    Window() : items_(N) {}
    inline void Push(T t)
    {
        items_[(size_t)((first_ + size_) % N)] = t;
        if (size_ < N) ++size_;
        else first_ = (first_ + 1) % N;
    }
    inline int GetSize() const { return size_; }
    inline T operator[](int index) const { return items_[(size_t)((first_ + index) % N)]; }

private:
    std::vector<T> items_;
    int first_ = 0, size_ = 0;
};

// the structure-of-arrays problem the C ABI takes (include/viso_ba.h: viso_ba_set_problem)
template <typename MapPointPtr> struct WindowProblem {
    std::vector<double> poses7;       // q (x, y, z, w), t — world -> camera (keyframe.h:82-89)
    std::vector<double> points3;
    std::vector<int32_t> edge_pose, edge_point;
    std::vector<float> edge_uv;       // cv::Point2f of the keypoint (keyframe.h:15)
    std::vector<MapPointPtr> points;  // who to write points3 back to
    int skipped_observations = 0;     // keypoint index outside Keypoints() (the latent bug at viso.cpp:88-92)
};

inline void QuaternionToRotation(const double q[4], double R[9] /* row-major */)
{
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    R[0] = 1 - 2 * (y * y + z * z), R[1] = 2 * (x * y - z * w), R[2] = 2 * (x * z + y * w);
    R[3] = 2 * (x * y + z * w), R[4] = 1 - 2 * (x * x + z * z), R[5] = 2 * (y * z - x * w);
    R[6] = 2 * (x * z - y * w), R[7] = 2 * (y * z + x * w), R[8] = 1 - 2 * (x * x + y * y);
}

// Map -> SoA.  `window`: the keyframes to adjust, oldest first (pose index = position; pose 0 is the gauge and stays
// fixed, as viso keeps its first frame at identity, keyframe.h:34-35).  A point enters with the observations that fall
// on window keyframes, if there are at least `min_observations` of them (one view does not constrain depth).
template <typename WindowT, typename MapPointPtr>
WindowProblem<MapPointPtr> ExtractWindow(WindowT& window, const std::vector<MapPointPtr>& map_points, int min_observations = 2)
{
    WindowProblem<MapPointPtr> out;
    const int nk = window.GetSize();
    std::map<const void*, int> pose_of;
    for (int k = 0; k < nk; ++k) {
        auto kf = window[k];
        pose_of[(const void*)&*kf] = k;
        auto R = kf->GetR();
        auto T = kf->GetT();
        double Rc[9], q[7];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) Rc[3 * c + r] = R(r, c); // column-major, Eigen's default
        BundleAdjuster::RotationToQuaternion(Rc, q);
        q[4] = T(0), q[5] = T(1), q[6] = T(2);
        out.poses7.insert(out.poses7.end(), q, q + 7);
    }
    for (const MapPointPtr& mp : map_points) {
        const auto& obs = mp->GetObservations(); // (Keyframe::Ptr, keypoint index), map_point.h:21-29
        int n_in = 0;
        for (const auto& ob : obs) {
            auto it = pose_of.find((const void*)&*ob.first);
            if (it == pose_of.end()) continue;
            if (ob.second < 0 || ob.second >= (int)ob.first->Keypoints().size()) continue;
            ++n_in;
        }
        if (n_in < min_observations) {
            for (const auto& ob : obs)
                if (pose_of.count((const void*)&*ob.first) && (ob.second < 0 || ob.second >= (int)ob.first->Keypoints().size())) ++out.skipped_observations;
            continue;
        }
        const int id = (int)out.points.size();
        auto P = mp->GetWorldPos();
        out.points3.push_back(P(0)), out.points3.push_back(P(1)), out.points3.push_back(P(2));
        out.points.push_back(mp);
        for (const auto& ob : obs) {
            auto it = pose_of.find((const void*)&*ob.first);
            if (it == pose_of.end()) continue;
            if (ob.second < 0 || ob.second >= (int)ob.first->Keypoints().size()) {
                ++out.skipped_observations;
                continue;
            }
            const auto& px = ob.first->Keypoints()[(size_t)ob.second].pt;
            out.edge_pose.push_back(it->second);
            out.edge_point.push_back(id);
            out.edge_uv.push_back((float)px.x), out.edge_uv.push_back((float)px.y);
        }
    }
    return out;
}

// One local bundle adjustment of the window: extract, optimise on the GPU, write back.  Types: Mat3 / Vec3 are what
// SetR / SetT / SetWorldPos take (M3d / V3d in viso).  Returns the C-ABI status; `summary` may be null.
template <typename Mat3, typename Vec3, typename WindowT, typename MapPointPtr>
int OptimizeWindow(BundleAdjuster& ba, WindowT& window, const std::vector<MapPointPtr>& map_points, double fx, double fy, double cx,
                   double cy, int iterations, viso_ba_summary* summary = nullptr, double huber_delta = 0.0)
{
    WindowProblem<MapPointPtr> pr = ExtractWindow(window, map_points);
    const int nk = window.GetSize();
    if (nk < 2 || pr.points.empty()) return VISO_BA_ERR_STATE;
    int st = ba.Clear();
    if (st == VISO_BA_OK) st = ba.SetIntrinsics(fx, fy, cx, cy);
    if (st == VISO_BA_OK) st = ba.SetHuberDelta(huber_delta);
    if (st != VISO_BA_OK) return st;
    st = viso_ba_set_problem(ba.Handle(), nk, pr.poses7.data(), (int64_t)pr.points.size(), pr.points3.data(), (int64_t)pr.edge_pose.size(),
                             pr.edge_pose.data(), pr.edge_point.data(), pr.edge_uv.data());
    if (st != VISO_BA_OK) return st;
    st = ba.Optimize(iterations, summary);
    if (st != VISO_BA_OK) return st;
    for (int k = 1; k < nk; ++k) { // pose 0 is the fixed gauge
        double q[7], R[9];
        if ((st = ba.GetPose(k, q)) != VISO_BA_OK) return st;
        QuaternionToRotation(q, R);
        Mat3 Rm;
        Vec3 Tm;
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) Rm(r, c) = R[3 * r + c];
            Tm(r) = q[4 + r];
        }
        auto kf = window[k];
        kf->SetR(Rm);
        kf->SetT(Tm);
    }
    for (size_t i = 0; i < pr.points.size(); ++i) {
        double p[3];
        if ((st = ba.GetPoint((int)i, p)) != VISO_BA_OK) return st;
        Vec3 Pm;
        for (int r = 0; r < 3; ++r) Pm(r) = p[r];
        pr.points[i]->SetWorldPos(Pm); // the setter MapPoint lacks (map_point.h:31): `inline void SetWorldPos(V3d P) { Pw_ = P; }`
    }
    return VISO_BA_OK;
}

} // namespace viso_ba_glue

// ---- inside viso (OpenCV + Eigen present): the two call sites, with the reference's own types ------------------------
#if defined(__has_include)
#if __has_include(<opencv2/core.hpp>) && __has_include(<Eigen/Core>) && defined(VISO_BA_GLUE_WITH_VISO_TYPES)
// viso.h gains:   BundleAdjuster ba_{20000, 32, 200000};   viso_ba_glue::Window<8, Keyframe::Ptr> window_;
// map_point.h gains:   inline void SetWorldPos(V3d P) { Pw_ = P; }
#include "map.h"
namespace viso_ba_glue {
// after the loop that ends at src/viso.cpp:96 (two-view initialisation), and after LKAlignment at src/viso.cpp:121 once the
// aligned positions kp_after[i] have been added to cur_frame as keypoints + observations of their map points:
template <int N> inline int AdjustLocalWindow(BundleAdjuster& ba, Window<N, Keyframe::Ptr>& window, Map& map, const M3d& K, int iterations = 10)
{
    std::vector<MapPoint::Ptr> points = map.GetPoints(); // by value, map.h:19
    return OptimizeWindow<M3d, V3d>(ba, window, points, K(0, 0), K(1, 1), K(0, 2), K(1, 2), iterations);
}
} // namespace viso_ba_glue
#endif
#endif

#endif // VISO_BA_GLUE_H
