"""CPU tests of the Lucas-Kanade oracle (oracle/lk_oracle.c) — what pins it, given that the reference
has no tests for src/viso.cpp:259-391 / :845-925: recovery of a known integer shift, recovery of the
true projections on a rendered plane scene, one Gauss-Newton step against an independent numpy
statement, and the reference's failure branches (border, flat patch)."""
import numpy as np
import pytest

import lk_oracle_lib as lko
import numpy_direct
from viso_b200.direct_synth import FX, FY, CX, CY, synth_scene, quat_to_R, pyramid


@pytest.fixture(scope="module")
def scene():
    sc = synth_scene(seed=7, n_points=200)
    P = sc["points"]
    sc["uv_ref"] = np.stack([FX * P[:, 0] / P[:, 2] + CX, FY * P[:, 1] / P[:, 2] + CY], axis=1)
    Pc = P @ quat_to_R(sc["true_pose"][:4]).T + sc["true_pose"][4:]
    sc["uv_true"] = np.stack([FX * Pc[:, 0] / Pc[:, 2] + CX, FY * Pc[:, 1] / Pc[:, 2] + CY], axis=1)
    return sc


@pytest.fixture(scope="module")
def shifted():
    sc = synth_scene(seed=9, n_points=10)
    ref = sc["ref_pyr"][0]
    cur = np.roll(ref, (-2, 3), axis=(0, 1))  # content moves by (+3, -2) pixels in (x, y)
    rng = np.random.default_rng(0)
    kp = np.stack([rng.uniform(60, 580, 64), rng.uniform(60, 420, 64)], axis=1)
    return pyramid(ref), pyramid(cur), kp, np.array([3.0, -2.0])


def test_alignment_recovers_an_integer_shift(shifted):
    ref_pyr, cur_pyr, kp, shift = shifted
    uv, ok = lko.align(lko.cfg(), [ref_pyr], 0, cur_pyr, kp, kp)
    assert ok.all()
    assert np.abs(uv - (kp + shift)).max() < 0.05


@pytest.mark.parametrize("inverse", [False, True])
def test_optical_flow_recovers_an_integer_shift(shifted, inverse):
    ref_pyr, cur_pyr, kp, shift = shifted
    kp2, ok = lko.flow(lko.cfg(), ref_pyr, cur_pyr, kp, None, inverse=inverse)
    assert kp2.dtype == np.float32 and ok.all()
    assert np.abs(kp2 - (kp + shift)).max() < 0.05
    # with an initial guess (kp2 non-empty, viso.cpp:267)
    kp2b, okb = lko.flow(lko.cfg(), ref_pyr, cur_pyr, kp, kp + shift + 0.4, inverse=inverse)
    assert okb.all() and np.abs(kp2b - (kp + shift)).max() < 0.05


def test_alignment_finds_the_true_projections_on_the_plane_scene(scene):
    uv, ok = lko.align(lko.cfg(), [scene["ref_pyr"]], 0, scene["cur_pyr"], scene["uv_ref"], scene["uv_ref"])
    assert ok.mean() > 0.95
    err0 = np.abs(scene["uv_ref"] - scene["uv_true"])[ok].max()
    err = np.abs(uv - scene["uv_true"])[ok]
    assert err0 > 3.0 and np.median(err) < 0.1 and err.max() < 0.5


def test_one_gauss_newton_step_matches_the_numpy_statement(scene):
    level, hp = 1, 4
    s = 0.5
    ref, cur = scene["ref_pyr"][level], scene["cur_pyr"][level]
    uv_ref, uv_cur = scene["uv_ref"][:20], scene["uv_ref"][:20] + 0.7
    out, ok = lko.align(lko.cfg(align_iterations=1), [scene["ref_pyr"]], 0, scene["cur_pyr"], uv_ref, uv_cur, level=level)
    d = np.arange(-hp, hp, dtype=np.float64)
    dx, dy = [a.ravel()[None, :] for a in np.meshgrid(d, d, indexing="ij")]
    U, V = uv_ref[:, :1] * s + dx, uv_ref[:, 1:] * s + dy
    gx = -0.5 * (numpy_direct.bilinear(ref, U + 1, V) - numpy_direct.bilinear(ref, U - 1, V))  # J = -grad(ref), viso.cpp:882-883
    gy = -0.5 * (numpy_direct.bilinear(ref, U, V + 1) - numpy_direct.bilinear(ref, U, V - 1))
    err = numpy_direct.bilinear(ref, U, V) - numpy_direct.bilinear(cur, uv_cur[:, :1] * s + dx, uv_cur[:, 1:] * s + dy)
    for i in range(len(uv_ref)):
        J = np.stack([gx[i], gy[i]], axis=1)
        upd = np.linalg.inv(J.T @ J) @ (-(J.T @ err[i]))
        assert np.allclose(out[i], uv_cur[i] + upd / s, rtol=0, atol=1e-9)  # viso.cpp:916


def test_failure_branches(scene):
    c = lko.cfg()
    # viso.cpp:869-873: the patch around uv_ref leaves the reference image -> success = false, no motion
    uv, ok = lko.align(c, [scene["ref_pyr"]], 0, scene["cur_pyr"], [[2.0, 2.0], [638.0, 478.0]], [[2.0, 2.0], [638.0, 478.0]])
    assert not ok.any() and np.array_equal(uv, [[2.0, 2.0], [638.0, 478.0]])
    # viso.cpp:896-899: a flat patch makes H singular -> NaN update -> success = false
    flat = pyramid(np.full((480, 640), 77, dtype=np.uint8))
    uv, ok = lko.align(c, [flat], 0, flat, [[300.0, 200.0]], [[301.0, 200.5]])
    assert not ok.any() and np.array_equal(uv, [[301.0, 200.5]])
    kp2, ok = lko.flow(c, flat, flat, [[300.0, 200.0]], None)
    assert not ok.any() and np.array_equal(kp2, np.array([[300.0, 200.0]], dtype=np.float32))
    # viso.cpp:286-289: keypoint within half_patch_size of the border
    kp2, ok = lko.flow(c, scene["ref_pyr"], scene["cur_pyr"], [[3.0, 100.0]], None, level=0)
    assert not ok.any()
