"""GPU parity of the Lucas-Kanade alignment (viso_b200/csrc/lk_align.cu) against the CPU oracle
(oracle/lk_oracle.c), through the C ABI.  Tolerance: both sides sum in fp64 but in a different fixed
order, and the reference's `cost > lastCost` stop (viso.cpp:900-902, :323-325) compares two nearly equal
numbers once the patch has converged, so the two sides may stop one tiny update apart: positions agree
to 1e-6 px (double path) / 1e-4 px (cv::Point2f path), success flags exactly."""
import numpy as np
import pytest

import lk_oracle_lib as lko
from viso_b200.direct_synth import FX, FY, CX, CY, synth_scene, quat_to_R, pyramid

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def scene():
    sc = synth_scene(seed=7, n_points=1500)
    P = sc["points"]
    sc["uv_ref"] = np.stack([FX * P[:, 0] / P[:, 2] + CX, FY * P[:, 1] / P[:, 2] + CY], axis=1)
    Pc = P @ quat_to_R(sc["true_pose"][:4]).T + sc["true_pose"][4:]
    sc["uv_true"] = np.stack([FX * Pc[:, 0] / Pc[:, 2] + CX, FY * Pc[:, 1] / Pc[:, 2] + CY], axis=1)
    sc2 = synth_scene(seed=8, n_points=10, rotvec=(-0.003, 0.004, 0.002), trans=(-0.01, 0.006, 0.008))
    sc["ref2_pyr"] = sc2["ref_pyr"]
    return sc


def _aligner(scene, **kw):
    from viso_b200 import lk
    a = lk.LKAligner(max_keypoints=2048, slots=3, **kw)
    a.set_pyramid(0, scene["ref_pyr"])
    a.set_pyramid(1, scene["cur_pyr"])
    a.set_pyramid(2, scene["ref2_pyr"])
    return a


def test_alignment_matches_the_oracle(vb, scene):
    a = _aligner(scene)
    uv, ok, ms = a.LKAlignment(0, 1, scene["uv_ref"], scene["uv_ref"])
    uvo, oko = lko.align(lko.cfg(), [scene["ref_pyr"]], 0, scene["cur_pyr"], scene["uv_ref"], scene["uv_ref"])
    assert np.array_equal(ok, oko) and ok.mean() > 0.95
    assert np.abs(uv - uvo).max() < 1e-6
    assert np.median(np.abs(uv - scene["uv_true"])[ok]) < 0.1
    assert ms > 0 and a.kernel_launches() == 1  # all four levels in one launch


def test_pairs_pick_their_own_reference_keyframe(vb, scene):
    a = _aligner(scene)
    n = 300
    rs = (np.arange(n) % 2) * 2  # slots 0 and 2 alternate (AlignmentPair::ref_frame, include/viso.h:116-121)
    uv, ok, _ = a.LKAlignment(rs, 1, scene["uv_ref"][:n], scene["uv_ref"][:n])
    uvo, oko = lko.align(lko.cfg(), [scene["ref_pyr"], scene["cur_pyr"], scene["ref2_pyr"]], rs, scene["cur_pyr"], scene["uv_ref"][:n],
                         scene["uv_ref"][:n])
    assert np.array_equal(ok, oko)
    assert np.abs(uv - uvo).max() < 1e-6


@pytest.mark.parametrize("level", [0, 2, 3])
def test_single_level_alignment_matches_the_oracle(vb, scene, level):
    a = _aligner(scene, align_iterations=3)
    start = scene["uv_ref"] + 0.8
    uv, ok = a.LKAlignmentSingle(level, 0, 1, scene["uv_ref"], start)
    uvo, oko = lko.align(lko.cfg(align_iterations=3), [scene["ref_pyr"]], 0, scene["cur_pyr"], scene["uv_ref"], start, level=level)
    assert np.array_equal(ok, oko)
    assert np.abs(uv - uvo).max() < 1e-8  # three forced iterations: no convergence-noise branch involved


@pytest.mark.parametrize("inverse", [False, True])
@pytest.mark.parametrize("initial", [False, True])
def test_optical_flow_matches_the_oracle(vb, scene, inverse, initial):
    a = _aligner(scene)
    kp1 = scene["uv_ref"].astype(np.float32)
    kp2 = (scene["uv_true"] + 0.6).astype(np.float32) if initial else None
    out, ok, ms = a.OpticalFlowMultiLevel(0, 1, kp1, kp2, inverse=inverse)
    outo, oko = lko.flow(lko.cfg(), scene["ref_pyr"], scene["cur_pyr"], kp1, kp2, inverse=inverse)
    assert out.dtype == np.float32 and np.array_equal(ok, oko) and ok.mean() > 0.9
    assert np.abs(out - outo).max() < 1e-4
    assert np.median(np.abs(out - scene["uv_true"])[ok]) < 0.1
    assert a.kernel_launches() == 1


def test_single_level_optical_flow_matches_the_oracle(vb, scene):
    a = _aligner(scene, flow_iterations=2)
    kp1 = (scene["uv_ref"] * 0.5).astype(np.float32)
    out, ok = a.OpticalFlowSingleLevel(1, 0, 1, kp1, None, inverse=False)
    outo, oko = lko.flow(lko.cfg(flow_iterations=2), scene["ref_pyr"], scene["cur_pyr"], kp1, None, inverse=False, level=1)
    assert np.array_equal(ok, oko)
    assert np.abs(out - outo).max() < 1e-5


def test_edge_cases(vb, scene):
    from viso_b200 import lk
    a = _aligner(scene)
    # empty input (LKAlignment with no visible map point)
    uv, ok, _ = a.LKAlignment(0, 1, np.zeros((0, 2)), np.zeros((0, 2)))
    assert uv.shape == (0, 2) and ok.shape == (0,)
    # border (viso.cpp:869-873, :286-289) and flat patch (viso.cpp:896-899, :315-321)
    pts = np.array([[2.0, 2.0], [638.0, 478.0], [320.0, 240.0]])
    uv, ok, _ = a.LKAlignment(0, 1, pts, pts)
    uvo, oko = lko.align(lko.cfg(), [scene["ref_pyr"]], 0, scene["cur_pyr"], pts, pts)
    assert list(ok) == list(oko) == [False, False, True] and np.array_equal(uv[:2], pts[:2])
    flat = pyramid(np.full((480, 640), 77, dtype=np.uint8))
    a.set_pyramid(2, flat)
    uv, ok, _ = a.LKAlignment(2, 2, [[300.0, 200.0]], [[301.0, 200.5]])
    assert not ok.any() and np.array_equal(uv, [[301.0, 200.5]])
    kp2, ok, _ = a.OpticalFlowMultiLevel(2, 2, [[300.0, 200.0]], None)
    assert not ok.any() and np.array_equal(kp2, np.array([[300.0, 200.0]], dtype=np.float32))
    # errors are status codes
    with pytest.raises(lk.LkError) as e:
        a.LKAlignment(7, 1, pts, pts)
    assert e.value.code == 2  # unknown slot
    with pytest.raises(lk.LkError) as e:
        a.LKAlignment(0, 1, np.zeros((5000, 2)) + 100, np.zeros((5000, 2)) + 100)
    assert e.value.code == 1  # capacity
    with pytest.raises(lk.LkError) as e:
        a.LKAlignment(0, 1, [[np.nan, 1.0]], [[1.0, 1.0]])
    assert e.value.code == 5
    fresh = lk.LKAligner()
    with pytest.raises(lk.LkError) as e:
        fresh.LKAlignment(0, 1, pts, pts)
    assert e.value.code == 7  # images not set


def test_runs_are_bit_identical(vb, scene):
    a = _aligner(scene)
    u1, o1, _ = a.LKAlignment(0, 1, scene["uv_ref"], scene["uv_ref"])
    u2, o2, _ = a.LKAlignment(0, 1, scene["uv_ref"], scene["uv_ref"])
    assert np.array_equal(u1, u2) and np.array_equal(o1, o2)


@pytest.mark.parametrize("hp", [2, 6, 8])
def test_other_patch_sizes_match_the_oracle(vb, scene, hp):
    """half_patch_size is a constant of Viso (include/viso.h:25) but a parameter here: 4x4 patches use one pixel
    per lane, 12x12 and 16x16 patches the general pixels-per-lane path of the kernels"""
    a = _aligner(scene, half_patch_size=hp, align_iterations=5, flow_iterations=4)
    n = 400
    start = scene["uv_ref"][:n] + 0.6
    uv, ok, _ = a.LKAlignment(0, 1, scene["uv_ref"][:n], start)
    c = lko.cfg(half_patch_size=hp, align_iterations=5, flow_iterations=4)
    uvo, oko = lko.align(c, [scene["ref_pyr"]], 0, scene["cur_pyr"], scene["uv_ref"][:n], start)
    assert np.array_equal(ok, oko)
    assert np.abs(uv - uvo).max() < 1e-7
    kp1 = scene["uv_ref"][:n].astype(np.float32)
    out, okf, _ = a.OpticalFlowMultiLevel(0, 1, kp1, None, inverse=False)
    outo, okfo = lko.flow(c, scene["ref_pyr"], scene["cur_pyr"], kp1, None, inverse=False)
    assert np.array_equal(okf, okfo)
    assert np.abs(out - outo).max() < 1e-4


def test_patch_size_beyond_the_kernel_limit_is_rejected(vb):
    from viso_b200 import lk
    with pytest.raises(lk.LkError) as e:
        lk.LKAligner(half_patch_size=9)  # 18 x 18 = 324 pixels > 32 lanes x 8
    assert e.value.code == 6
