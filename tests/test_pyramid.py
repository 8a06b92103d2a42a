"""Image pyramid (Keyframe's constructor, reference include/keyframe.h:29-45 -> cv::pyrDown).
This is the one piece of the path whose reference arithmetic is PINNED: tests/golden/pyrdown_cv2.npz
holds outputs of OpenCV's own cv2.pyrDown (tools/make_golden_pyr.py), the library the reference calls.
CPU: the numpy restatement reproduces them bit for bit.  GPU: so does the device kernel, through
the C ABI, and the pyramids it builds feed the tracker kernels."""
import os

import numpy as np
import pytest

import numpy_pyr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = np.load(os.path.join(ROOT, "tests", "golden", "pyrdown_cv2.npz"))
CASES = sorted({k[:-3] for k in GOLDEN.files if k.endswith("_l0")})


def test_golden_file_has_the_expected_cases():
    assert CASES == ["extremes_48x64", "rand_120x160", "rand_37x100", "rand_61x81", "texture_480x640"]
    assert GOLDEN["texture_480x640_l3"].shape == (60, 80)   # 640x480 -> 320x240 -> 160x120 -> 80x60
    assert GOLDEN["rand_61x81_l1"].shape == (30, 40)        # Size(cols * 0.5, rows * 0.5) truncates odd sizes
    assert GOLDEN["rand_37x100_l3"].shape == (4, 12)


@pytest.mark.parametrize("case", CASES)
def test_numpy_statement_matches_opencv_bit_for_bit(case):
    pyr = numpy_pyr.pyramid(GOLDEN[case + "_l0"])
    for level in range(4):
        assert np.array_equal(pyr[level], GOLDEN["%s_l%d" % (case, level)]), (case, level)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_device_pyramid_matches_opencv_bit_for_bit(vb, case):
    from viso_b200 import lk
    a = lk.LKAligner(slots=2)
    a.set_frame(1, GOLDEN[case + "_l0"])
    assert a.kernel_launches() == 3
    for level in range(4):
        got = a.get_image(1, level)
        want = GOLDEN["%s_l%d" % (case, level)]
        assert got.shape == want.shape and np.array_equal(got, want), (case, level)


@pytest.mark.gpu
def test_device_pyramid_of_a_strided_image(vb):
    """cv::Mat rows may be padded (step > cols): the row stride is honoured"""
    from viso_b200 import lk
    img = GOLDEN["rand_61x81_l0"]
    padded = np.zeros((61, 96), dtype=np.uint8)
    padded[:, :81] = img
    a = lk.LKAligner(slots=1)
    a.set_frame(0, padded[:, :81])  # a view: strides[0] = 96
    for level in range(4):
        assert np.array_equal(a.get_image(0, level), GOLDEN["rand_61x81_l%d" % level])


@pytest.mark.gpu
def test_too_small_an_image_is_an_error(vb):
    from viso_b200 import lk
    a = lk.LKAligner(slots=1)
    with pytest.raises(lk.LkError) as e:
        a.set_frame(0, np.zeros((7, 9), dtype=np.uint8))  # 9x7 -> 4x3 -> 2x1 -> 1x0
    assert e.value.code == 6


@pytest.mark.gpu
def test_tracker_kernels_on_device_built_pyramids_match_the_oracle(vb):
    """set_frame (level 0 only) + device pyrDown gives the tracker kernels the same pyramids OpenCV would:
    direct pose estimation and LK alignment then agree with their oracles run on the golden pyramids"""
    import direct_oracle_lib as dorc
    import lk_oracle_lib as lko
    from viso_b200 import direct, lk
    from viso_b200.direct_synth import FX, FY, CX, CY, synth_scene
    sc = synth_scene(seed=4, n_points=600)  # its reference image is the golden texture_480x640
    ref_pyr = [GOLDEN["texture_480x640_l%d" % l] for l in range(4)]
    assert np.array_equal(sc["ref_pyr"][0], ref_pyr[0])
    cur_pyr = numpy_pyr.pyramid(sc["cur_pyr"][0])
    est = direct.DirectPoseEstimator(max_points=600, reference_termination=0)
    est.set_frame(0, ref_pyr[0])
    est.set_frame(1, cur_pyr[0])
    est.set_points(sc["points"])
    pose, stats, _ = est.DirectPoseEstimationMultiLayer(sc["identity"])
    po, so = dorc.multi_layer(dorc.cfg(reference_termination=0), ref_pyr, cur_pyr, sc["ref_pose"], sc["identity"], sc["points"])
    assert [s.iterations for s in stats[0]] == [s.iterations for s in so]
    assert np.abs(pose - po).max() < 1e-9
    P = sc["points"]
    uv_ref = np.stack([FX * P[:, 0] / P[:, 2] + CX, FY * P[:, 1] / P[:, 2] + CY], axis=1)
    a = lk.LKAligner(slots=2)
    a.set_frame(0, ref_pyr[0])
    a.set_frame(1, cur_pyr[0])
    uv, ok, _ = a.LKAlignment(0, 1, uv_ref, uv_ref)
    uvo, oko = lko.align(lko.cfg(), [ref_pyr], 0, cur_pyr, uv_ref, uv_ref)
    assert np.array_equal(ok, oko) and np.abs(uv - uvo).max() < 1e-6
