"""EdgeDirectProjection::computeError (reference include/bundle_adjuster.h:70-101): the
reference adjuster's only edge, as a batched kernel.  CPU: the oracle against a hand-written numpy statement
with the reference's float casts; GPU: the kernel against the oracle through the C ABI."""
import numpy as np
import pytest

import lk_oracle_lib as lko
import numpy_direct
from viso_b200.direct_synth import FX, FY, CX, CY, synth_scene, quat_to_R

K = (FX, FY, CX, CY)


@pytest.fixture(scope="module")
def scene():
    sc = synth_scene(seed=13, n_points=800)
    # a few points whose patch leaves an image: near the left / top border and far outside
    extra = np.array([[(1.2 - CX) / FX, 0.0, 1.0], [0.0, (0.4 - CY) / FY, 1.0], [3.0, 0.0, 1.0], [(-0.5 - CX) / FX, (100 - CY) / FY, 1.0]])
    sc["pts"] = np.vstack([sc["points"], extra])
    sc["frame_poses"] = np.stack([sc["ref_pose"], sc["true_pose"]])
    sc["vertex_poses"] = np.stack([sc["ref_pose"], sc["true_pose"]])
    return sc


def _numpy_errors(sc, pts, src, tgt):
    """bundle_adjuster.h:74-99, one edge at a time, float casts as at :86-89"""
    out = np.zeros((len(pts), 16))
    lvl = np.zeros(len(pts), dtype=np.uint8)
    pyr = [sc["ref_pyr"], sc["cur_pyr"]]
    for e, p in enumerate(pts):
        T = sc["vertex_poses"][tgt]
        Pc = quat_to_R(T[:4]) @ p + T[4:]
        tu, tv = (FX * Pc[0] + CX * Pc[2]) / Pc[2], (FY * Pc[1] + CY * Pc[2]) / Pc[2]
        F = sc["frame_poses"][src]
        Ps = quat_to_R(F[:4]) @ p + F[4:]
        su, sv = Ps[0] / Ps[2] * FX + CX, Ps[1] / Ps[2] * FY + CY
        idx = 0
        for j in range(-2, 2):
            stop = False
            for i in range(-2, 2):
                u1, v1 = np.float32(tu + i), np.float32(tv + j)
                u2, v2 = np.float32(su + i), np.float32(sv + j)
                a, b = pyr[tgt][0], pyr[src][0]
                if not (0 <= u1 < a.shape[1] and 0 <= v1 < a.shape[0] and 0 <= u2 < b.shape[1] and 0 <= v2 < b.shape[0]):
                    lvl[e] = 1
                    stop = True
                    break
                out[e, idx] = (numpy_direct.bilinear(b, np.array([float(u2)]), np.array([float(v2)]))[0]
                               - numpy_direct.bilinear(a, np.array([float(u1)]), np.array([float(v1)]))[0])
                idx += 1
            if stop:
                break
    return out, lvl


def test_oracle_matches_the_numpy_statement(scene):
    pts = scene["pts"][-60:]
    err, lvl = lko.photo_edge_errors([scene["ref_pyr"], scene["cur_pyr"]], pts, 0, 1, scene["frame_poses"], scene["vertex_poses"], K)
    want, wl = _numpy_errors(scene, pts, 0, 1)
    assert np.array_equal(lvl, wl) and list(lvl[-4:]) == [1, 1, 1, 1] and lvl[:-4].sum() == 0
    assert np.allclose(err, want, rtol=0, atol=1e-10)
    # the first sample (j = -2, i = -2) of the top-border point fails at once; the left-border point fails at i = -2 too
    assert not err[-3].any()


def test_same_frame_same_pose_gives_zero_error_and_true_motion_small_error(scene):
    pyr = [scene["ref_pyr"], scene["cur_pyr"]]
    err, lvl = lko.photo_edge_errors(pyr, scene["points"], 0, 0, scene["frame_poses"], scene["vertex_poses"], K)
    assert not lvl.any() and np.abs(err).max() < 1e-9  # float casts of equal doubles are equal: identical samples
    err, lvl = lko.photo_edge_errors(pyr, scene["points"], 0, 1, scene["frame_poses"], scene["vertex_poses"], K)
    wrong = np.stack([scene["ref_pose"], scene["ref_pose"]])
    err_wrong, _ = lko.photo_edge_errors(pyr, scene["points"], 0, 1, scene["frame_poses"], wrong, K)
    assert np.abs(err).mean() < 0.25 * np.abs(err_wrong).mean()  # the true pose explains the image


@pytest.mark.gpu
def test_kernel_matches_the_oracle(vb, scene):
    from viso_b200 import lk
    a = lk.LKAligner(max_keypoints=1024, slots=2)
    a.set_pyramid(0, scene["ref_pyr"])
    a.set_pyramid(1, scene["cur_pyr"])
    n = len(scene["pts"])
    src = np.arange(n) % 2
    tgt = 1 - src
    err, lvl, ms = a.EdgeDirectProjection_computeError(scene["pts"], src, tgt, scene["frame_poses"], scene["vertex_poses"], K)
    eo, lo = lko.photo_edge_errors([scene["ref_pyr"], scene["cur_pyr"]], scene["pts"], src, tgt, scene["frame_poses"], scene["vertex_poses"], K)
    assert np.array_equal(lvl, lo) and lvl.sum() >= 4
    assert np.abs(err - eo).max() < 1e-10
    assert ms > 0 and a.kernel_launches() == 1
    # an odd number of edges (a half-filled warp) and the empty set
    e1, l1, _ = a.EdgeDirectProjection_computeError(scene["pts"][:7], 0, 1, scene["frame_poses"], scene["vertex_poses"], K)
    eo1, lo1 = lko.photo_edge_errors([scene["ref_pyr"], scene["cur_pyr"]], scene["pts"][:7], 0, 1, scene["frame_poses"], scene["vertex_poses"], K)
    assert np.array_equal(l1, lo1) and np.abs(e1 - eo1).max() < 1e-10
    e0, l0, _ = a.EdgeDirectProjection_computeError(np.zeros((0, 3)), 0, 1, scene["frame_poses"], scene["vertex_poses"], K)
    assert e0.shape == (0, 16) and l0.shape == (0,)
