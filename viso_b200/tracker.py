"""The tracking branch of Viso::OnNewFrame (reference src/viso.cpp:112-137) on the B200 kernels, plus the
local bundle adjustment the reference declares but never calls (call-site patch (ii) in
INTEGRATION.md).  Host logic in Python because the reference's own C++ needs OpenCV / Eigen / Sophus, which
this image lacks; every numeric kernel is a C-ABI call into libviso_ba.so (no CPU fallback):

    DirectPoseEstimationMultiLayer(cur_frame, X)     viso.cpp:114-115  -> viso_direct_estimate
    LKAlignment(cur_frame, kp_before, kp_after)      viso.cpp:120-121  -> viso_lk_align (+ the pair selection
                                                                          of viso.cpp:772-826 below, host side)
    BundleAdjuster over the last N keyframes         bundle_adjuster.h -> viso_ba_* (window path)
"""
import numpy as np

from . import api as _api
from . import direct as _direct
from . import lk as _lk
from .direct_synth import FX, FY, CX, CY, quat_to_R


class Tracker:
    def __init__(self, max_points=4096, max_keyframes=8, width=640, height=480, device=0, K=(FX, FY, CX, CY),
                 reference_termination=1):
        self.K = K
        self.w, self.h = width, height
        self.max_keyframes = max_keyframes
        self.direct = _direct.DirectPoseEstimator(max_points=max_points, device=device, width=width, height=height,
                                                  fx=K[0], fy=K[1], cx=K[2], cy=K[3], reference_termination=reference_termination)
        self.lk = _lk.LKAligner(max_keypoints=max_points, slots=max_keyframes + 1, device=device, width=width, height=height)
        self.device = device
        self.keyframes = []   # (image slot, pose [7]) — Map::Keyframes()
        self.points = None    # Map::GetPoints() world positions [n][3]
        self.last = None      # (image, pose) of last_frame

    # -- Map ---------------------------------------------------------------------------------------
    def add_keyframe(self, image, pose):
        slot = len(self.keyframes)
        if slot >= self.max_keyframes:
            raise ValueError("keyframe slots exhausted")
        self.lk.set_frame(slot, image)  # Keyframe(cv::Mat): pyramid built on the device
        self.keyframes.append((slot, np.array(pose, dtype=np.float64)))
        self.last = (image, np.array(pose, dtype=np.float64))
        return slot

    def set_map_points(self, xyz):
        self.points = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        self.direct.set_points(self.points)

    def _project(self, pose, P):
        Pc = P @ quat_to_R(pose[:4]).T + pose[4:]
        fx, fy, cx, cy = self.K
        return np.stack([fx * Pc[:, 0] / Pc[:, 2] + cx, fy * Pc[:, 1] / Pc[:, 2] + cy], axis=1), Pc

    def _inside(self, uv):
        return (uv[:, 0] >= 0) & (uv[:, 0] < self.w) & (uv[:, 1] >= 0) & (uv[:, 1] < self.h)

    # -- viso.cpp:772-826: for every map point visible in the current frame, the keyframe with the best viewing angle
    def select_alignment_pairs(self, cur_pose, max_angle=180.0):
        P = self.points
        uv_cur, _ = self._project(cur_pose, P)
        visible = self._inside(uv_cur)                         # current_frame->IsInside(Pw, 0), viso.cpp:777-779
        best_angle = np.full(len(P), 180.0)                    # viso.cpp:782
        best_idx = np.full(len(P), -1, dtype=np.int32)
        best_uv = np.zeros((len(P), 2))
        for j, (slot, pose) in enumerate(self.keyframes):      # viso.cpp:787-808
            uv_ref, Pc = self._project(pose, P)
            ok = self._inside(uv_ref)
            n = Pc / np.linalg.norm(Pc, axis=1, keepdims=True)
            angle = np.abs(np.arccos(np.clip(n[:, 2], -1, 1)) / np.pi * 180)   # Keyframe::ViewingAngle, keyframe.h:93-98
            take = ok & ~((angle > max_angle) | (angle > best_angle))
            best_angle[take], best_idx[take] = angle[take], j
            best_uv[take] = uv_ref[take]
        sel = visible & (best_idx >= 0)                        # viso.cpp:810-812
        return np.nonzero(sel)[0], best_idx[sel], best_uv[sel], uv_cur[sel]

    # -- viso.cpp:112-137 -----------------------------------------------------------------------------
    def track(self, image):
        """One frame of the kRunning branch.  Returns dict(pose, point_index, kp_before, kp_after, ms_direct, ms_lk)."""
        last_image, last_pose = self.last
        self.direct.set_frame(0, last_image)
        self.direct.set_frame(1, image)
        self.direct.set_reference_pose(last_pose)
        X, _, ms_d = self.direct.DirectPoseEstimationMultiLayer(last_pose)   # X starts at last_frame's pose, viso.cpp:113-115
        cur_slot = self.max_keyframes
        self.lk.set_frame(cur_slot, image)
        idx, ref_kf, uv_ref, uv_cur = self.select_alignment_pairs(X)
        ref_slot = np.array([self.keyframes[j][0] for j in ref_kf], dtype=np.int32)
        aligned, ok, ms_l = self.lk.LKAlignment(ref_slot, cur_slot, uv_ref, uv_cur)
        self.last = (image, X)
        return {"pose": X, "point_index": idx[ok], "kp_before": uv_cur[ok], "kp_after": aligned[ok],   # viso.cpp:832-843, :919-923
                "ref_keyframe": ref_kf[ok], "uv_ref": uv_ref[ok], "ms_direct": ms_d, "ms_lk": ms_l}

    # -- the local bundle adjustment the reference declares (bundle_adjuster.h) but never calls ----------
    def bundle_adjust(self, window_poses, observations, iterations=10, huber_delta=2.0):
        """window_poses [N][7] (pose 0 stays fixed); observations: list over the window's frames of
        (point_index [m], uv [m][2]).  Returns (refined poses, refined points, Summary)."""
        used = np.unique(np.concatenate([o[0] for o in observations]))
        remap = -np.ones(len(self.points), dtype=np.int64)
        remap[used] = np.arange(len(used))
        edge_pose = np.concatenate([np.full(len(o[0]), k, dtype=np.int32) for k, o in enumerate(observations)])
        edge_point = np.concatenate([remap[o[0]] for o in observations]).astype(np.int32)
        edge_uv = np.concatenate([o[1] for o in observations]).astype(np.float32)
        ba = _api.BundleAdjuster(len(used), len(window_poses), len(edge_pose), device=self.device, huber_delta=huber_delta)
        ba.set_problem(np.asarray(window_poses, dtype=np.float64), self.points[used], edge_pose, edge_point, edge_uv)
        summary = ba.Optimize(iterations)
        poses, pts = ba.poses().copy(), ba.points().copy()
        ba.close()
        refined = self.points.copy()
        refined[used] = pts
        return poses, refined, summary
