"""The CPU oracle of photometric bundle adjustment (oracle/photo_ba_oracle.c): what pins it.
The reference holds no vectors for this path and cannot run (bundle_adjuster.h:154-165 does not compile), so the
checks are: an independent numpy statement of the residual and its analytic Jacobian, central differences of the
smooth residual (the analytic Jacobian uses central-difference image gradients, so the agreement is approximate by
construction), and recovery of a perturbed synthetic window."""
import numpy as np
import pytest

import numpy_photo
import oracle_lib as orc
import photo_ba_oracle_lib as porc
from viso_b200.direct_synth import synth_photo_window


@pytest.fixture(scope="module")
def win():
    return synth_photo_window(seed=3, n_frames=4, n_points=120)


def test_edge_residual_and_jacobian_match_the_numpy_statement(win):
    c = orc.cfg(K=win["K"])
    n_valid = 0
    for e in range(0, len(win["edge_point"]), 7):
        s, d, p = win["edge_src"][e], win["edge_dst"][e], win["edge_point"][e]
        args = (win["images"][s], win["images"][d], win["poses_init"][s], win["poses_init"][d], win["points_init"][p])
        ok, r, Jp, Jl = porc.edge_jacobian(c, *args)
        ok2, r2, Jp2, Jl2 = numpy_photo.edge(win["K"], *args)
        assert ok == ok2
        n_valid += ok
        np.testing.assert_allclose(r, r2, atol=1e-9)
        np.testing.assert_allclose(Jp, Jp2, rtol=1e-10, atol=1e-9)
        np.testing.assert_allclose(Jl, Jl2, rtol=1e-10, atol=1e-9)
    assert n_valid > 30


def test_analytic_jacobian_follows_central_differences_of_the_smooth_residual(win):
    """d r / d point and d r / d pose by central differences of the residual without the float casts.  The analytic form
    uses the reference's central-difference image gradient on 8-bit images, so the two agree to a few per cent."""
    from viso_b200.direct_synth import pose_from_rotvec, quat_to_R
    K = win["K"]
    num, den = 0.0, 0.0
    for e in range(0, len(win["edge_point"]), 23):
        s, d, p = win["edge_src"][e], win["edge_dst"][e], win["edge_point"][e]
        si, ti, sp, dp, P = win["images"][s], win["images"][d], win["poses_init"][s], win["poses_init"][d], win["points_init"][p]
        ok, r, Jp, Jl = numpy_photo.edge(K, si, ti, sp, dp, P, smooth=True)
        if not ok:
            continue
        h = 2e-3  # about one pixel of motion: the sampler is piecewise bilinear, smaller steps see its kinks
        Jl_fd = np.zeros((16, 3))
        for a in range(3):
            dP = np.zeros(3)
            dP[a] = h
            Jl_fd[:, a] = (numpy_photo.edge(K, si, ti, sp, dp, P + dP, True)[1] - numpy_photo.edge(K, si, ti, sp, dp, P - dP, True)[1]) / (2 * h)
        Jp_fd = np.zeros((16, 6))
        R, t = quat_to_R(dp[:4]), dp[4:]
        for a in range(6):
            rs = []
            for sgn in (+1, -1):
                xi = np.zeros(6)
                xi[a] = sgn * h
                ex = pose_from_rotvec(xi[3:], xi[:3])  # exp for a pure translation or a pure rotation
                Rn = quat_to_R(ex[:4]) @ R
                tn = quat_to_R(ex[:4]) @ t + ex[4:]
                # back to a quaternion pose: reuse numpy_photo through an explicit pose vector
                qw = 0.5 * np.sqrt(max(1e-30, 1 + np.trace(Rn)))
                q = np.array([(Rn[2, 1] - Rn[1, 2]) / (4 * qw), (Rn[0, 2] - Rn[2, 0]) / (4 * qw), (Rn[1, 0] - Rn[0, 1]) / (4 * qw), qw])
                rs.append(numpy_photo.edge(K, si, ti, sp, np.concatenate([q, tn]), P, True)[1])
            Jp_fd[:, a] = (rs[0] - rs[1]) / (2 * h)
        num += np.linalg.norm(Jl - Jl_fd)**2 + np.linalg.norm(Jp - Jp_fd)**2
        den += np.linalg.norm(Jl_fd)**2 + np.linalg.norm(Jp_fd)**2
    assert den > 0 and np.sqrt(num / den) < 0.15, np.sqrt(num / den)


def test_lm_reduces_the_photometric_cost_and_pulls_the_window_towards_the_truth(win):
    c = orc.cfg(K=win["K"])
    c0 = porc.cost(c, win["images"], win["poses_init"], win["poses_init"], win["points_init"], win["edge_point"], win["edge_src"], win["edge_dst"])
    fixed = [1, 1] + [0] * (len(win["poses_init"]) - 2)  # keyframes 0 and 1 pin the gauge including the scale
    poses, points, fc, it, stats = porc.optimize(c, win["images"], win["poses_init"], win["points_init"], win["edge_point"],
                                                 win["edge_src"], win["edge_dst"], 10, fixed=fixed)
    assert it >= 1 and abs(stats[0].cost_before - c0) <= 1e-9 * c0
    assert fc < 0.1 * c0, (c0, fc)
    assert all(s.cost_after <= s.cost_before for s in stats)
    e0 = np.abs(win["poses_init"][2:] - win["poses_true"][2:]).max()
    e1 = np.abs(poses[2:] - win["poses_true"][2:]).max()
    assert e1 < 0.7 * e0, (e0, e1)
    np.testing.assert_array_equal(poses[:2], win["poses_init"][:2])


def test_an_edge_that_leaves_the_image_contributes_nothing(win):
    c = orc.cfg(K=win["K"])
    pts = win["points_init"].copy()
    pts[0] = [5.0, 5.0, 1.0]  # projects far outside every image
    sel = win["edge_point"] == 0
    full = porc.cost(c, win["images"], win["poses_init"], win["poses_init"], pts, win["edge_point"], win["edge_src"], win["edge_dst"])
    rest = porc.cost(c, win["images"], win["poses_init"], win["poses_init"], pts, win["edge_point"][~sel], win["edge_src"][~sel], win["edge_dst"][~sel])
    assert full == rest
