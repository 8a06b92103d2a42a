"""CPU tests of the direct-pose oracle (oracle/direct_oracle.c) — what pins it, given that the
reference has no tests for src/viso.cpp:661-766: known answers of the bilinear sampler and the
gradient, dPixeldXi entry by entry against the formula in the reference text and against finite
differences of Project(exp(eps) T, P), an independent numpy statement of the normal equations,
the reference's termination quirk, and recovery of a known camera motion."""
import ctypes as C

import numpy as np
import pytest

import direct_oracle_lib as dorc
import numpy_direct
import oracle_lib as orc
from viso_b200.direct_synth import FX, FY, CX, CY, synth_scene, pose_from_rotvec, quat_to_R

K = (FX, FY, CX, CY)


@pytest.fixture(scope="module")
def scene():
    return synth_scene(seed=3, n_points=300)


def test_pixel_value_known_answers():
    img = np.arange(48, dtype=np.uint8).reshape(6, 8) * 3
    assert dorc.pixel_value(img, 2.0, 3.0) == float(img[3, 2])
    # keyframe.h:52-56: weights (1-xx)(1-yy), xx(1-yy), (1-xx)yy, xx yy on data[0], [1], [step], [step+1]
    want = 0.75 * 0.5 * img[1, 4] + 0.25 * 0.5 * img[1, 5] + 0.75 * 0.5 * img[2, 4] + 0.25 * 0.5 * img[2, 5]
    assert dorc.pixel_value(img, 4.25, 1.5) == pytest.approx(want, abs=1e-12)
    # the last row: data[step] leaves the buffer in the reference; the oracle clamps the address
    assert np.isfinite(dorc.pixel_value(img, 7.5, 5.5))
    g = dorc.gradient(img, 3.0, 2.0)  # keyframe.h:59-64: central differences of the sampler
    assert g[0] == pytest.approx(0.5 * (float(img[2, 4]) - float(img[2, 2])))
    assert g[1] == pytest.approx(0.5 * (float(img[3, 3]) - float(img[1, 3])))


def test_dpixel_dxi_matches_the_reference_formula_and_finite_differences():
    c = dorc.cfg()
    rng = np.random.default_rng(0)
    for scale in (1.0, 0.5, 0.125):
        pose = pose_from_rotvec(rng.normal(0, 0.05, 3), rng.normal(0, 0.05, 3))
        P = np.array([0.2, -0.1, 1.3])
        R, T = quat_to_R(pose[:4]), pose[4:]
        J = np.zeros(12)
        dorc.lib().direct_dpixel_dxi(C.byref(c), R.ravel().ctypes.data_as(dorc._D), np.ascontiguousarray(T).ctypes.data_as(dorc._D),
                                     P.ctypes.data_as(dorc._D), scale, J.ctypes.data_as(dorc._D))
        x, y, z = R @ P + T
        fx, fy = FX * scale, FY * scale
        want = [fx / z, 0, -fx * x / z**2, -fx * x * y / z**2, fx + fx * x * x / z**2, -fx * y / z,
                0, fy / z, -fy * y / z**2, -fy - fy * y * y / z**2, fy * x * y / z**2, fy * x / z]  # viso.cpp:653-655
        assert np.allclose(J, want, rtol=1e-14, atol=0)
        # d Project(exp(eps) T, P) / d eps, central differences
        num = np.zeros((2, 6))
        for k in range(6):
            uv = []
            for s in (+1e-6, -1e-6):
                xi = np.zeros(6)
                xi[k] = s
                p2 = pose.copy()
                orc.lib().oracle_pose_oplus(xi.ctypes.data_as(dorc._D), p2.ctypes.data_as(dorc._D))
                out = np.zeros(2)
                dorc.lib().direct_project(C.byref(c), p2.ctypes.data_as(dorc._D), P.ctypes.data_as(dorc._D), scale, out.ctypes.data_as(dorc._D))
                uv.append(out)
            num[:, k] = (uv[0] - uv[1]) / 2e-6
        assert np.allclose(J.reshape(2, 6), num, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("level", [0, 1, 3])
def test_normal_equations_match_the_numpy_statement(scene, level):
    c = dorc.cfg()
    T21 = pose_from_rotvec((0.003, -0.004, 0.002), (0.01, -0.006, 0.008))
    H, b, cost, ng = dorc.linearize(c, level, scene["ref_pyr"][level], scene["cur_pyr"][level], scene["ref_pose"], T21, scene["points"])
    H2, b2, cost2, ng2 = numpy_direct.linearize(K, level, scene["ref_pyr"][level], scene["cur_pyr"][level], scene["ref_pose"], T21,
                                                scene["points"])
    assert ng == ng2 and ng > 0
    assert np.allclose(H, H2, rtol=1e-10, atol=1e-6 * np.abs(H2).max())
    assert np.allclose(b, b2, rtol=1e-10, atol=1e-9 * np.abs(b2).max())
    assert cost == pytest.approx(cost2, rel=1e-12)
    assert np.allclose(H, H.T, rtol=0, atol=1e-9 * np.abs(H).max())


def test_points_outside_the_image_are_skipped(scene):
    c = dorc.cfg()
    pts = np.vstack([scene["points"][:50], [[5.0, 0.0, 1.0], [-0.625, 0.0, 1.0], [0.0, 0.0, -1.0]]])
    args = (c, 0, scene["ref_pyr"][0], scene["cur_pyr"][0], scene["ref_pose"], scene["identity"])
    _, _, _, ng50 = dorc.linearize(*args, pts[:50])
    _, _, _, ng52 = dorc.linearize(*args, pts[:52])
    _, _, _, ng53 = dorc.linearize(*args, pts)
    assert ng52 == ng50  # viso.cpp:704-715: a patch that leaves either image is skipped
    assert ng53 == ng50 + 1  # ... but a point BEHIND the camera projects to (cx, cy) and is kept: no depth test there


def test_reference_loop_stops_after_one_step_per_level(scene):
    """viso.cpp:673, :751-753: lastCost starts at 0, so (1 - cost/lastCost) = -inf < delta_thresh on
    iteration 0 and the loop breaks after ONE Gauss-Newton update.  Restated, not repaired."""
    c = dorc.cfg()
    pose, st = dorc.multi_layer(c, scene["ref_pyr"], scene["cur_pyr"], scene["ref_pose"], scene["identity"], scene["points"])
    assert [s.iterations for s in st] == [1, 1, 1, 1]
    assert [s.stop_reason for s in st] == [3, 3, 3, 3]
    assert not np.allclose(pose, scene["identity"])


def test_repaired_loop_recovers_the_camera_motion(scene):
    c = dorc.cfg(reference_termination=0)
    pose, st = dorc.multi_layer(c, scene["ref_pyr"], scene["cur_pyr"], scene["ref_pose"], scene["identity"], scene["points"])
    assert sum(s.iterations for s in st) > 4
    err0 = np.abs(scene["identity"] - scene["true_pose"]).max()
    err = np.abs(pose - scene["true_pose"]).max()
    assert err < 0.1 * err0, (err, err0)
    assert err < 1.5e-3
