"""GPU parity of the direct pose estimator (viso_b200/csrc/direct_pose.cu) against the CPU oracle
(oracle/direct_oracle.c), through the C ABI.  Tolerances: the sums are fp64 on both sides but in a
different (fixed) order and with FMA contraction on the device, so H, b and the cost agree to
1e-10 relative, poses to 1e-9."""
import numpy as np
import pytest

import direct_oracle_lib as dorc
from viso_b200.direct_synth import synth_scene, pose_from_rotvec

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def scene():
    return synth_scene(seed=5, n_points=1500)


def _estimator(vb, sc, frames=1, **kw):
    from viso_b200 import direct
    est = direct.DirectPoseEstimator(max_points=len(sc["points"]), frames=frames, **kw)
    for f in range(frames):
        est.set_pyramid(0, sc["ref_pyr"], frame=f)
        est.set_pyramid(1, sc["cur_pyr"], frame=f)
        est.set_reference_pose(sc["ref_pose"], frame=f)
    est.set_points(sc["points"])
    return est


@pytest.mark.parametrize("level", [0, 1, 2, 3])
def test_normal_equations_match_the_oracle(vb, scene, level):
    est = _estimator(vb, scene)
    T21 = pose_from_rotvec((0.002, -0.003, 0.001), (0.008, -0.004, 0.006))
    H, b, cost, ng = est.linearize(level, T21)
    Ho, bo, co, ngo = dorc.linearize(dorc.cfg(), level, scene["ref_pyr"][level], scene["cur_pyr"][level], scene["ref_pose"], T21,
                                     scene["points"])
    assert ng[0] == ngo and ngo > 1000
    assert np.allclose(H[0], Ho, rtol=1e-10, atol=1e-11 * np.abs(Ho).max())
    assert np.allclose(b[0], bo, rtol=1e-10, atol=1e-11 * np.abs(bo).max())
    assert cost[0] == pytest.approx(co, rel=1e-12)
    assert est.kernel_launches() == 1


def test_reference_loop_matches_the_oracle(vb, scene):
    """the loop exactly as the reference writes it: one Gauss-Newton step per level"""
    est = _estimator(vb, scene)
    pose, stats, ms = est.DirectPoseEstimationMultiLayer(scene["identity"])
    po, so = dorc.multi_layer(dorc.cfg(), scene["ref_pyr"], scene["cur_pyr"], scene["ref_pose"], scene["identity"], scene["points"])
    assert np.abs(pose - po).max() < 1e-9
    for l in range(4):
        assert (stats[0][l].iterations, stats[0][l].n_good, stats[0][l].stop_reason) == (so[l].iterations, so[l].n_good, so[l].stop_reason)
        assert stats[0][l].cost == pytest.approx(so[l].cost, rel=1e-9)
    assert [s.iterations for s in stats[0]] == [1, 1, 1, 1]
    assert ms > 0 and est.kernel_launches() == 1  # the whole multi-level loop is one launch


def test_repaired_loop_matches_the_oracle_and_recovers_the_motion(vb, scene):
    est = _estimator(vb, scene, reference_termination=0)
    pose, stats, _ = est.DirectPoseEstimationMultiLayer(scene["identity"])
    po, so = dorc.multi_layer(dorc.cfg(reference_termination=0), scene["ref_pyr"], scene["cur_pyr"], scene["ref_pose"],
                              scene["identity"], scene["points"])
    assert [s.iterations for s in stats[0]] == [s.iterations for s in so]
    assert [s.stop_reason for s in stats[0]] == [s.stop_reason for s in so]
    assert np.abs(pose - po).max() < 1e-9
    assert np.abs(pose - scene["true_pose"]).max() < 1.5e-3


def test_single_layer_matches_the_oracle(vb, scene):
    est = _estimator(vb, scene, reference_termination=0, max_iterations=7)
    T21 = pose_from_rotvec((0.001, 0.0, 0.0), (0.004, 0.0, 0.003))
    pose, st = est.DirectPoseEstimationSingleLayer(1, T21)
    po, so = dorc.single_layer(dorc.cfg(reference_termination=0, max_iterations=7), 1, scene["ref_pyr"][1], scene["cur_pyr"][1],
                               scene["ref_pose"], T21, scene["points"])
    assert (st[0].iterations, st[0].stop_reason, st[0].n_good) == (so.iterations, so.stop_reason, so.n_good)
    assert np.abs(pose - po).max() < 1e-9


def test_batched_frames_are_independent_and_deterministic(vb, scene):
    frames = 5
    est = _estimator(vb, scene, frames=frames, reference_termination=0)
    rng = np.random.default_rng(2)
    init = np.stack([pose_from_rotvec(rng.normal(0, 0.002, 3), rng.normal(0, 0.004, 3)) for _ in range(frames)])
    p1, s1, _ = est.DirectPoseEstimationMultiLayer(init)
    p2, _, _ = est.DirectPoseEstimationMultiLayer(init)
    assert np.array_equal(p1, p2)  # fixed-shape sums: bit-identical from run to run
    single = _estimator(vb, scene, reference_termination=0)
    for f in range(frames):
        pf, _, _ = single.DirectPoseEstimationMultiLayer(init[f])
        assert np.array_equal(pf, p1[f])  # a frame's result does not depend on its neighbours in the batch
        po, _ = dorc.multi_layer(dorc.cfg(reference_termination=0), scene["ref_pyr"], scene["cur_pyr"], scene["ref_pose"], init[f],
                                 scene["points"])
        assert np.abs(pf - po).max() < 1e-9


def test_edge_cases(vb, scene):
    from viso_b200 import direct
    est = _estimator(vb, scene)
    # every point outside the image: nGood = 0, H singular -> NaN update -> best pose restored (viso.cpp:741-744)
    est.set_points(np.array([[9.0, 9.0, 1.0], [-9.0, 2.0, 1.0]]))
    pose, st = est.DirectPoseEstimationSingleLayer(0, scene["identity"])
    po, so = dorc.single_layer(dorc.cfg(), 0, scene["ref_pyr"][0], scene["cur_pyr"][0], scene["ref_pose"], scene["identity"],
                               np.array([[9.0, 9.0, 1.0], [-9.0, 2.0, 1.0]]))
    assert (st[0].n_good, st[0].stop_reason) == (so.n_good, so.stop_reason) == (0, 1)
    assert np.array_equal(pose, scene["identity"]) and np.array_equal(po, scene["identity"])
    # capacity, state and argument errors are status codes, not crashes
    with pytest.raises(direct.DirectError) as e:
        est.set_points(np.zeros((len(scene["points"]) + 1, 3)) + 1.0)
    assert e.value.code == 1
    with pytest.raises(direct.DirectError) as e:
        est.set_points(np.array([[np.nan, 0, 1.0]]))
    assert e.value.code == 5
    fresh = direct.DirectPoseEstimator()
    fresh.set_points(scene["points"][:10])
    with pytest.raises(direct.DirectError) as e:
        fresh.DirectPoseEstimationMultiLayer(scene["identity"])
    assert e.value.code == 7  # images not set


def test_moved_reference_frame_strided_images_and_small_patch(vb, scene):
    """last_frame away from the origin (the map's world frame is not the last frame's), cv::Mat rows with padding
    (step > cols), and a 4x4 patch"""
    from viso_b200 import direct
    from viso_b200.direct_synth import quat_to_R
    ref_pose = pose_from_rotvec((0.01, -0.02, 0.015), (0.05, -0.03, 0.02))
    R, t = quat_to_R(ref_pose[:4]), ref_pose[4:]
    world_pts = (scene["points"] - t) @ R          # Pc = R Pw + t  ->  Pw = R^T (Pc - t)
    # T21 maps the world to the current frame: (true motion) o (reference pose)
    est = direct.DirectPoseEstimator(max_points=len(world_pts), reference_termination=0, half_patch_size=2, max_iterations=6)

    def padded(img):
        buf = np.zeros((img.shape[0], img.shape[1] + 24), dtype=np.uint8)
        buf[:, :img.shape[1]] = img
        return buf[:, :img.shape[1]]  # a view whose row stride is cols + 24
    ref_p, cur_p = [padded(im) for im in scene["ref_pyr"]], [padded(im) for im in scene["cur_pyr"]]
    assert ref_p[0].strides[0] == ref_p[0].shape[1] + 24
    est.set_pyramid(0, ref_p)
    est.set_pyramid(1, cur_p)
    est.set_reference_pose(ref_pose)
    est.set_points(world_pts)
    pose, stats, _ = est.DirectPoseEstimationMultiLayer(ref_pose)  # X starts at last_frame's pose (viso.cpp:113)
    c = dorc.cfg(reference_termination=0, half_patch_size=2, max_iterations=6)
    po, so = dorc.multi_layer(c, ref_p, cur_p, ref_pose, ref_pose, world_pts)
    assert [s.iterations for s in stats[0]] == [s.iterations for s in so]
    assert [s.n_good for s in stats[0]] == [s.n_good for s in so]
    assert np.abs(pose - po).max() < 1e-9
    # the recovered pose composes the true motion with the reference pose
    Rt, tt = quat_to_R(scene["true_pose"][:4]), scene["true_pose"][4:]
    want_R, want_t = Rt @ R, Rt @ t + tt
    assert np.abs(quat_to_R(pose[:4]) - want_R).max() < 3e-3 and np.abs(pose[4:] - want_t).max() < 3e-3
