"""the tracking branch of Viso::OnNewFrame (reference src/viso.cpp:112-137) followed by the
local bundle adjustment of INTEGRATION.md's call-site patch (ii), chained on the B200 kernels through
the three C ABIs (viso_b200/tracker.py).  A synthetic camera trajectory over a textured plane with known
poses: the direct estimator recovers every frame's pose from the previous one, LK alignment refines the
projections to sub-pixel observations, and bundle adjustment over the window of keyframes reduces the
error the noisy map put into poses and points."""
import numpy as np
import pytest

from viso_b200.direct_synth import synth_sequence, project

pytestmark = pytest.mark.gpu


def test_direct_pose_then_lk_alignment_then_bundle_adjustment(vb):
    from viso_b200.tracker import Tracker
    seq = synth_sequence(seed=21, n_frames=5, n_points=1200, step_rot=(0.002, -0.003, 0.002), step_trans=(0.03, -0.02, 0.01))
    rng = np.random.default_rng(5)
    true_pts = seq["points"]
    # the map as two-view initialisation left it: 1 % depth error along the rays of keyframe 0 (which sits at the origin)
    noisy_pts = true_pts * (1.0 + rng.normal(0, 0.01, (len(true_pts), 1)))
    tr = Tracker(max_points=len(true_pts), max_keyframes=6, reference_termination=0)
    tr.add_keyframe(seq["images"][0], seq["poses"][0])
    tr.set_map_points(noisy_pts)
    # keyframe 0's observations: the keypoints the map was triangulated from (cv::Point2f)
    obs = [(np.arange(len(true_pts)), (project(seq["poses"][0], true_pts) + rng.normal(0, 0.1, (len(true_pts), 2))).astype(np.float32))]
    tracked = [seq["poses"][0]]
    for k in range(1, 5):
        out = tr.track(seq["images"][k])
        err = np.abs(out["pose"] - seq["poses"][k]).max()
        assert err < 6e-3, (k, err)                                   # direct pose estimation from the previous frame
        truth = project(seq["poses"][k], true_pts)[out["point_index"]]
        assert len(out["point_index"]) > 0.9 * len(true_pts)
        assert np.median(np.abs(out["kp_after"] - truth)) < 0.25      # LK alignment lands on the true projections ...
        assert np.median(np.abs(out["kp_before"] - truth)) > np.median(np.abs(out["kp_after"] - truth))  # ... from worse ones
        assert out["ms_direct"] > 0 and out["ms_lk"] > 0
        obs.append((out["point_index"], out["kp_after"].astype(np.float32)))
        tracked.append(out["pose"])
        tr.add_keyframe(seq["images"][k], out["pose"])
    tracked = np.stack(tracked)
    poses, pts, summary = tr.bundle_adjust(tracked, obs, iterations=10, huber_delta=2.0)
    assert summary.final_cost < 0.1 * summary.initial_cost  # measured: 652 -> 28.6
    assert np.array_equal(poses[0], tracked[0])                        # gauge: pose 0 stays fixed
    # monocular BA leaves the scale free: compare after a least-squares scale fit of the structure
    def scale(p):
        return float((p * true_pts).sum() / (p * p).sum())
    s0, s1 = scale(noisy_pts), scale(pts)
    pt_err_before, pt_err_after = np.abs(noisy_pts * s0 - true_pts).mean(), np.abs(pts * s1 - true_pts).mean()
    t_err_before = np.abs(tracked[:, 4:] * s0 - seq["poses"][:, 4:]).max()
    t_err_after = np.abs(poses[:, 4:] * s1 - seq["poses"][:, 4:]).max()
    q_err_before, q_err_after = np.abs(tracked[:, :4] - seq["poses"][:, :4]).max(), np.abs(poses[:, :4] - seq["poses"][:, :4]).max()
    assert pt_err_after < 0.6 * pt_err_before, (pt_err_before, pt_err_after)   # measured: 3.7e-3 -> 1.8e-3
    assert t_err_after < 0.8 * t_err_before, (t_err_before, t_err_after)       # measured: 2.7e-3 -> 1.5e-3
    assert q_err_after < 0.8 * q_err_before, (q_err_before, q_err_after)       # measured: 1.25e-3 -> 0.72e-3
