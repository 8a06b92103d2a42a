"""the photometric edge of the reference adjuster (reference include/bundle_adjuster.h:58-106)
as a residual of the LM loop on the device, through the C ABI, against oracle/photo_ba_oracle.c (which forms one dense
system with explicit 16-row Jacobians and no Schur complement).  Tolerances: BASELINE.json north_star (final cost 1e-5
relative, poses 1e-6, points 1e-5); the reduced system and the cost are compared much tighter."""
import numpy as np
import pytest

import oracle_lib as orc
import photo_ba_oracle_lib as porc
import viso_b200 as vb
from viso_b200.direct_synth import synth_photo_window

pytestmark = pytest.mark.gpu


def _handle(win, fixed=(0, 1), **cfg):
    K = win["K"]
    ba = vb.BundleAdjuster(len(win["points_init"]), len(win["poses_init"]), len(win["edge_point"]), device=0,
                           fx=K[0], fy=K[1], cx=K[2], cy=K[3], **cfg)
    ba.set_photo_problem(win["poses_init"], win["points_init"], win["edge_point"], win["edge_src"], win["edge_dst"], win["images"])
    for f in fixed:
        ba.SetPoseFixed(f, True)
    return ba


def _fixed_mask(win, fixed):
    m = np.zeros(len(win["poses_init"]), np.uint8)
    m[list(fixed)] = 1
    return m


@pytest.fixture(scope="module")
def win():
    return synth_photo_window(seed=5, n_frames=4, n_points=150)


def test_cost_of_the_start_state_matches_the_oracle(win):
    ba = _handle(win)
    c = orc.cfg(K=win["K"])
    want = porc.cost(c, win["images"], win["poses_init"], win["poses_init"], win["points_init"], win["edge_point"], win["edge_src"], win["edge_dst"])
    got = ba.eval_cost()
    assert abs(got - want) <= 1e-9 * want, (got, want)


@pytest.mark.parametrize("huber", [0.0, 60.0])
def test_lm_iterations_match_the_oracle(win, huber):
    fixed = (0, 1)
    ba = _handle(win, fixed, huber_delta=huber)
    s, stats = ba.Optimize(6, with_stats=True)
    c = orc.cfg(K=win["K"], huber_delta=huber)
    poses, points, fc, it, ostats = porc.optimize(c, win["images"], win["poses_init"], win["points_init"], win["edge_point"],
                                                  win["edge_src"], win["edge_dst"], 6, fixed=_fixed_mask(win, fixed))
    assert s.iterations == it
    assert [x.trials for x in stats] == [x.trials for x in ostats]
    assert abs(s.initial_cost - ostats[0].cost_before) <= 1e-9 * ostats[0].cost_before
    assert s.final_cost < 0.2 * s.initial_cost
    assert abs(s.final_cost - fc) <= 1e-5 * fc, (s.final_cost, fc)
    assert np.abs(ba.poses() - poses).max() < 1e-6
    assert np.abs(ba.points() - points).max() < 1e-5
    names = set()
    ba.set_kernel_timing(True)
    ba.reset_state()
    ba.optimize_resident(1)
    names = set(ba.kernel_times())
    assert {"photo_sample", "photo_linearize", "photo_cost", "schur_blocks", "update"} <= names, names


def test_incremental_interface_and_edges_that_leave_the_image(win):
    """AddPoint / AddPose / AddPhotoEdge give the same result as the bulk call; a point that projects outside every image
    disables its edges (bundle_adjuster.h:91-95) on both sides."""
    pts = win["points_init"].copy()
    pts[3] = [4.0, 3.0, 1.0]
    K = win["K"]
    ba = vb.BundleAdjuster(len(pts), len(win["poses_init"]), len(win["edge_point"]), device=0, fx=K[0], fy=K[1], cx=K[2], cy=K[3])
    for q in win["poses_init"]:
        ba.AddPose(q)
    for p in pts:
        ba.AddPoint(p)
    for k, im in enumerate(win["images"]):
        ba.SetPoseImage(k, im)
    for l, s_, d in zip(win["edge_point"], win["edge_src"], win["edge_dst"]):
        ba.AddPhotoEdge(l, s_, d)
    ba.SetPoseFixed(0, True), ba.SetPoseFixed(1, True)
    s = ba.Optimize(3)
    c = orc.cfg(K=K)
    poses, points, fc, it, _ = porc.optimize(c, win["images"], win["poses_init"], pts, win["edge_point"], win["edge_src"], win["edge_dst"], 3,
                                             fixed=_fixed_mask(win, (0, 1)))
    assert s.iterations == it and abs(s.final_cost - fc) <= 1e-5 * fc
    assert np.abs(ba.poses() - poses).max() < 1e-6 and np.abs(ba.points() - points).max() < 1e-5
    np.testing.assert_array_equal(ba.points()[3], pts[3])  # no edge ever saw it: it does not move
    with pytest.raises(vb.VisoBaError):
        ba.AddEdge(0, 1, 10.0, 10.0)  # reprojection and photometric edges do not mix


def test_a_pose_without_an_image_is_refused(win):
    K = win["K"]
    ba = vb.BundleAdjuster(len(win["points_init"]), len(win["poses_init"]), len(win["edge_point"]), device=0, fx=K[0], fy=K[1], cx=K[2], cy=K[3])
    ba.set_photo_problem(win["poses_init"], win["points_init"], win["edge_point"], win["edge_src"], win["edge_dst"], win["images"][:-1])
    with pytest.raises(vb.VisoBaError):
        ba.Optimize(1)
