// bundle_adjuster.h — source-compatible replacement of viso's include/bundle_adjuster.h.
//
// The reference class (reference include/bundle_adjuster.h:23-187) is a g2o wrapper whose three
// methods have empty bodies (:169-173) and that never optimises or reads back.  This header
// keeps the class name, the default constructor (:111) and the three method names and
// arities (AddPoint :169, AddPose :171, AddEdge :173), and completes the interface with
// measurement-carrying edges, Optimize and getters.  Every call forwards to the C ABI of
// viso_ba.h (libviso_ba.so: hand-written sm_100a kernels; there is no CPU fallback).
//
// It compiles with no Eigen / Sophus / g2o / OpenCV.  When <Eigen/Core> and <sophus/se3.hpp>
// are on the include path the reference-typed overloads (V3d, Sophus::SE3d) are enabled too.
#ifndef VISO_BUNDLE_ADJUSTER_H
#define VISO_BUNDLE_ADJUSTER_H

#include "viso_ba.h"

#include <stdexcept>
#include <string>
#include <vector>

#if defined(__has_include)
#if __has_include(<Eigen/Core>) && __has_include(<sophus/se3.hpp>) && !defined(VISO_BA_NO_EIGEN)
#include <Eigen/Core>
#include <sophus/se3.hpp>
#define VISO_BA_HAVE_SOPHUS 1
#endif
#endif

class BundleAdjuster {
public:
    // bundle_adjuster.h:111 + :182-184: the pools are allocated once; 1000 each by default
    explicit BundleAdjuster(int max_points = 1000, int max_poses = 1000, int max_edges = 1000, int device = 0)
    {
        viso_ba_config cfg;
        viso_ba_default_config(&cfg);
        cfg.max_points = max_points;
        cfg.max_poses = max_poses;
        cfg.max_edges = max_edges;
        cfg.device = device;
        const int32_t st = viso_ba_create(&cfg, &h_);
        if (st != VISO_BA_OK) throw std::runtime_error(std::string("viso_ba_create: ") + viso_ba_status_string(st));
    }
    ~BundleAdjuster() { viso_ba_destroy(h_); }
    BundleAdjuster(const BundleAdjuster&) = delete;
    BundleAdjuster& operator=(const BundleAdjuster&) = delete;

    // ---- the reference's three methods (raw-pointer twins; return the new index, < 0 = -status)
    inline int AddPoint(const double xyz[3]) { return (int)viso_ba_add_point(h_, xyz); }
    // Sophus SE3d::data() order: unit quaternion (x, y, z, w) then translation; world -> camera
    inline int AddPose(const double q_xyzw_t[7]) { return (int)viso_ba_add_pose(h_, q_xyzw_t); }
    // Eigen default storage: R column-major
    inline int AddPose(const double R_colmajor[9], const double t[3])
    {
        double q[7];
        RotationToQuaternion(R_colmajor, q);
        q[4] = t[0], q[5] = t[1], q[6] = t[2];
        return AddPose(q);
    }
    // bundle_adjuster.h:173 takes (point, pose) only because its photometric edge reads the
    // images; the reprojection edge needs the measured pixel (Keypoints()[idx].pt, float32)
    inline void AddEdge(int point, int pose, double u, double v) { last_ = viso_ba_add_edge(h_, point, pose, (float)u, (float)v); }
    // reference signature, kept so that existing call sites still compile: the measurement is
    // supplied later with SetEdgeMeasurement; an edge without one is rejected by Optimize
    inline void AddEdge(int point, int pose)
    {
        pending_.push_back({ point, pose });
    }
    // The reference's own edge (bundle_adjuster.h:58-106, :154-160): EdgeDirectProjection(srcFrame, targetFrame) between point
    // and pose.  The keyframe of a pose vertex is its level-0 image (cv::Mat CV_8UC1: data, rows, cols, step); the source
    // keyframe is named by its pose index and projects with the pose it had when Optimize was called (:79).
    inline void SetPoseImage(int pose, const unsigned char* data, int rows, int cols, int step) { last_ = viso_ba_set_pose_image(h_, pose, data, rows, cols, step); }
    inline void AddPhotoEdge(int point, int src_pose, int dst_pose) { last_ = viso_ba_add_photo_edge(h_, point, src_pose, dst_pose); }
    inline void SetEdgeMeasurement(int pending_index, double u, double v)
    {
        if (pending_index < 0 || pending_index >= (int)pending_.size()) {
            last_ = VISO_BA_ERR_INDEX;
            return;
        }
        AddEdge(pending_[(size_t)pending_index].point, pending_[(size_t)pending_index].pose, u, v);
        pending_[(size_t)pending_index].point = -1;
    }

#ifdef VISO_BA_HAVE_SOPHUS
    inline int AddPoint(Eigen::Vector3d p) { return AddPoint(p.data()); }          // :169
    inline int AddPose(Sophus::SE3d p) { return AddPose(p.data()); }               // :171
    inline Sophus::SE3d GetPose(int i)
    {
        double q[7];
        GetPose(i, q);
        return Sophus::SE3d(Eigen::Quaterniond(q[3], q[0], q[1], q[2]), Eigen::Vector3d(q[4], q[5], q[6]));
    }
    inline Eigen::Vector3d GetPoint(int i)
    {
        Eigen::Vector3d p;
        GetPoint(i, p.data());
        return p;
    }
#endif

    // ---- what the reference never had -------------------------------------------------------
    inline int SetIntrinsics(double fx, double fy, double cx, double cy) { return last_ = viso_ba_set_intrinsics(h_, fx, fy, cx, cy); }
    inline int SetPoseFixed(int pose, bool fixed = true) { return last_ = viso_ba_set_pose_fixed(h_, pose, fixed ? 1 : 0); }
    inline int SetHuberDelta(double delta) { return last_ = viso_ba_set_huber_delta(h_, delta); }
    // optimizer_.initializeOptimization(); optimizer_.optimize(iterations);  returns a status
    inline int Optimize(int iterations, viso_ba_summary* summary = nullptr, viso_ba_iter_stats* stats = nullptr)
    {
        for (const Pending& e : pending_)
            if (e.point >= 0) return last_ = VISO_BA_ERR_STATE; // an AddEdge(point, pose) never got its pixel
        return last_ = viso_ba_optimize(h_, iterations, stats, summary);
    }
    inline int GetPose(int i, double out_q_xyzw_t[7]) { return last_ = viso_ba_get_poses(h_, i, 1, out_q_xyzw_t); }
    inline int GetPoint(int i, double out_xyz[3]) { return last_ = viso_ba_get_points(h_, i, 1, out_xyz); }
    inline int Clear()
    {
        pending_.clear();
        return last_ = viso_ba_clear(h_);
    }
    inline int NumPoints() const { return (int)viso_ba_num_points(h_); }
    inline int NumPoses() const { return (int)viso_ba_num_poses(h_); }
    inline int NumEdges() const { return (int)viso_ba_num_edges(h_); }
    inline int LastStatus() const { return last_; }
    inline const char* LastError() const { return viso_ba_last_error(h_); }
    inline viso_ba* Handle() { return h_; }

    static inline void RotationToQuaternion(const double Rc[9], double q[4])
    { // R column-major: R(r, c) = Rc[3 * c + r]
        const double r00 = Rc[0], r10 = Rc[1], r20 = Rc[2], r01 = Rc[3], r11 = Rc[4], r21 = Rc[5], r02 = Rc[6], r12 = Rc[7], r22 = Rc[8];
        const double tr = r00 + r11 + r22;
        double x, y, z, w;
        if (tr > 0.0) {
            const double s = 2.0 * Sqrt(tr + 1.0);
            w = 0.25 * s, x = (r21 - r12) / s, y = (r02 - r20) / s, z = (r10 - r01) / s;
        } else if (r00 > r11 && r00 > r22) {
            const double s = 2.0 * Sqrt(1.0 + r00 - r11 - r22);
            w = (r21 - r12) / s, x = 0.25 * s, y = (r01 + r10) / s, z = (r02 + r20) / s;
        } else if (r11 > r22) {
            const double s = 2.0 * Sqrt(1.0 + r11 - r00 - r22);
            w = (r02 - r20) / s, x = (r01 + r10) / s, y = 0.25 * s, z = (r12 + r21) / s;
        } else {
            const double s = 2.0 * Sqrt(1.0 + r22 - r00 - r11);
            w = (r10 - r01) / s, x = (r02 + r20) / s, y = (r12 + r21) / s, z = 0.25 * s;
        }
        q[0] = x, q[1] = y, q[2] = z, q[3] = w;
    }

private:
    static inline double Sqrt(double a)
    { // Newton on 1/sqrt would need <cmath>; keep the header's includes minimal but correct
        return __builtin_sqrt(a);
    }
    struct Pending {
        int point, pose;
    };
    viso_ba* h_ = nullptr;
    int last_ = VISO_BA_OK;
    std::vector<Pending> pending_;
};

#endif // VISO_BUNDLE_ADJUSTER_H
