"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/viso_ba.h declares, and the staging calls keep the error behaviour of the header
(status codes instead of the reference's std::cerr + exit(0), bundle_adjuster.h:133-135,
:162-165).  No compute call is made here (no GPU)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "viso_ba.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(viso_ba_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(vb):
    declared = _declared_symbols()
    assert len(declared) >= 35
    lib = C.CDLL(vb.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), "libviso_ba.so does not export " + name
    assert sorted(vb.SYMBOLS) == declared  # the Python binding covers the whole header
    out = subprocess.run(["nm", "-D", "--defined-only", vb.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (viso_ba_[a-z0-9_]+)", out))
    assert set(declared) <= exported


def test_synth_library_is_host_only_and_exports_its_header(vb):
    """include/viso_ba_synth.h lives in libviso_synth.so, which links no CUDA: the CPU reference arm of bench.py
    generates its problems without ever mapping the CUDA backend"""
    src = open(os.path.join(ROOT, "include", "viso_ba_synth.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    declared = sorted(set(re.findall(r"\b(viso_ba_synth_[a-z0-9_]+)\s*\(", src)))
    assert sorted(vb.SYNTH_SYMBOLS) == declared and len(declared) == 5
    out = subprocess.run(["nm", "-D", "--defined-only", vb.SYNTH_LIB_PATH], capture_output=True, text=True).stdout
    assert set(declared) <= set(re.findall(r" T (viso_ba_synth_[a-z0-9_]+)", out))
    needed = subprocess.run(["readelf", "-d", vb.SYNTH_LIB_PATH], capture_output=True, text=True).stdout
    assert "cuda" not in needed.lower() and "nccl" not in needed.lower()
    # a fresh interpreter that only generates a problem maps libviso_synth.so and not libviso_ba.so
    code = ("import sys; sys.path.insert(0, %r); import viso_b200 as vb; p = vb.synth_config(1, n_points=50, target_edges=200);"
            "m = open('/proc/self/maps').read(); print(int('libviso_synth.so' in m), int('libviso_ba.so' in m))" % ROOT)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, check=True)
    assert res.stdout.split() == ["1", "0"]


def test_version_config_and_status_strings(vb):
    lib = vb.load_library()
    assert lib.viso_ba_version() == 100
    cfg = vb.default_config()
    assert cfg.struct_size == C.sizeof(vb.Config)
    assert (cfg.max_points, cfg.max_poses, cfg.max_edges) == (1000, 1000, 1000)  # bundle_adjuster.h:182-184
    assert (cfg.fx, cfg.fy, cfg.cx, cfg.cy) == (517.3, 516.5, 325.1, 249.7)  # src/main.cpp:14-17
    assert cfg.huber_delta == 0.0 and cfg.lm_tau == 1e-5 and cfg.lm_max_trials == 10
    assert lib.viso_ba_status_string(0) == b"ok"
    assert lib.viso_ba_status_string(3) == b"CUDA error"
    bad = vb.Config()
    bad.struct_size = 12
    h = C.c_void_p()
    assert lib.viso_ba_create(C.byref(bad), C.byref(h)) == 6  # ERR_ARG


def test_reference_named_methods_and_pool_capacity(vb):
    ba = vb.BundleAdjuster()  # reference default: pools of 1000
    assert ba.AddPoint([0.0, 0.0, 1.0]) == 0  # returns the new index (bundle_adjuster.h:169)
    assert ba.AddPoint([0.1, 0.0, 1.0]) == 1
    assert ba.AddPose([0, 0, 0, 1, 0, 0, 0]) == 0  # bundle_adjuster.h:171
    ba.AddEdge(1, 0, 320.0, 240.0)  # (point, pose) order of bundle_adjuster.h:173
    assert (ba.num_points, ba.num_poses, ba.num_edges) == (2, 1, 1)
    ba.Clear()
    assert (ba.num_points, ba.num_poses, ba.num_edges) == (0, 0, 0)
    small = vb.BundleAdjuster(max_points=2, max_poses=1, max_edges=1)
    small.AddPoint([0, 0, 1])
    small.AddPoint([0, 0, 1])
    with pytest.raises(vb.VisoBaError) as e:
        small.AddPoint([0, 0, 1])
    assert e.value.status == 1  # ERR_CAPACITY instead of the reference's cerr + continue
    small.AddPose([0, 0, 0, 1, 0, 0, 0])
    with pytest.raises(vb.VisoBaError) as e:
        small.AddPose([0, 0, 0, 1, 0, 0, 0])
    assert e.value.status == 1
    small.AddEdge(0, 0, 1, 1)
    with pytest.raises(vb.VisoBaError) as e:
        small.AddEdge(1, 0, 1, 1)
    assert e.value.status == 1  # the reference exit(0)s here (bundle_adjuster.h:162-165)


def test_bad_indices_and_non_finite_inputs(vb):
    ba = vb.BundleAdjuster()
    ba.AddPoint([0, 0, 1])
    ba.AddPose([0, 0, 0, 1, 0, 0, 0])
    for args in ((1, 0), (0, 1), (-1, 0)):
        with pytest.raises(vb.VisoBaError) as e:
            ba.AddEdge(args[0], args[1], 1.0, 1.0)
        assert e.value.status == 2  # ERR_INDEX
    with pytest.raises(vb.VisoBaError) as e:
        ba.AddPoint([np.nan, 0, 1])
    assert e.value.status == 5
    with pytest.raises(vb.VisoBaError) as e:
        ba.AddPose([0, 0, 0, 0, 0, 0, 0])  # zero quaternion
    assert e.value.status == 6
    with pytest.raises(vb.VisoBaError) as e:
        ba.AddEdge(0, 0, float("inf"), 1.0)
    assert e.value.status == 5
    with pytest.raises(vb.VisoBaError) as e:
        ba.SetPoseFixed(5)
    assert e.value.status == 2
    big = vb.BundleAdjuster(10, 10, 10)
    with pytest.raises(vb.VisoBaError) as e:
        big.set_problem(np.tile([0, 0, 0, 1, 0, 0, 0.0], (2, 1)), np.ones((3, 3)), [0, 2], [0, 1], np.ones((2, 2)))
    assert e.value.status == 2
    with pytest.raises(vb.VisoBaError) as e:
        big.set_problem(np.tile([0, 0, 0, 1, 0, 0, 0.0], (2, 1)), np.ones((30, 3)), [0], [0], np.ones((1, 2)))
    assert e.value.status == 1
    with pytest.raises(vb.VisoBaError) as e:  # batched: indices are local to the problem
        big.set_problem_batched([0, 1, 2], [0, 1, 2], [0, 1, 2], np.tile([0, 0, 0, 1, 0, 0, 0.0], (2, 1)),
                                np.ones((2, 3)), [0, 1], [0, 0], np.ones((2, 2)))
    assert e.value.status == 2


def test_get_before_optimize_returns_the_staged_values(vb):
    ba = vb.BundleAdjuster()
    ba.AddPoint([1.0, 2.0, 3.0])
    ba.AddPose([0, 0, 0, 2.0, 4, 5, 6])  # stored as given; normalised on upload
    assert np.array_equal(ba.GetPoint(0), [1, 2, 3])
    assert np.array_equal(ba.GetPose(0), [0, 0, 0, 2, 4, 5, 6])
    with pytest.raises(vb.VisoBaError):
        ba.GetPoint(1)


def test_compute_without_a_gpu_fails_loudly(vb):
    """no CPU fallback: on a box without a usable device every compute entry point errors"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    ba = vb.BundleAdjuster()
    ba.AddPoint([0, 0, 1])
    ba.AddPose([0, 0, 0, 1, 0, 0, 0])
    ba.AddEdge(0, 0, 300.0, 200.0)
    for call in (lambda: ba.Optimize(1), ba.upload, ba.eval_cost, lambda: ba.optimize_resident(1)):
        with pytest.raises(vb.VisoBaError) as e:
            call()
        assert e.value.status == 3  # ERR_CUDA
        assert "no CPU fallback" in str(e.value)


def test_missing_library_is_an_import_error(vb, tmp_path):
    with pytest.raises(ImportError):
        vb.load_library(str(tmp_path / "libviso_ba.so"))


def test_synthetic_generator_is_deterministic_and_well_formed(vb):
    a, b = vb.synth_config(1), vb.synth_config(1)
    for k in ("poses_init", "points_init", "edge_pose", "edge_point", "edge_uv"):
        assert np.array_equal(a[k], b[k])
    assert a["poses_init"].shape == (10, 7) and a["points_init"].shape == (2000, 3)
    assert abs(len(a["edge_pose"]) - 15000) < 600
    assert a["edge_uv"].dtype == np.float32  # cv::Point2f, keyframe.h:15
    assert np.array_equal(a["poses_init"][0], [0, 0, 0, 1, 0, 0, 0])  # frame 0 = identity (keyframe.h:34-35)
    assert np.bincount(a["edge_point"], minlength=2000).min() >= 2
    assert np.bincount(a["edge_pose"], minlength=10).min() >= 1
    uv = a["edge_uv"]
    assert uv[:, 0].min() > -10 and uv[:, 0].max() < 650 and uv[:, 1].min() > -10 and uv[:, 1].max() < 490
    c2 = vb.synth_config(2)
    assert c2["huber_delta"] == 2.0 and abs(len(c2["edge_pose"]) - 80000) < 3000
    w0, w1 = vb.synth_config(5, 0), vb.synth_config(5, 1)
    assert not np.array_equal(w0["points_init"], w1["points_init"])
    assert abs(a["points_true"][:, 2].mean() - 1.0) < 0.1  # mean depth 1 (viso.cpp:622-637)


def test_host_structure_build_does_not_depend_on_the_thread_count():
    """the general path's host build (sort by point and pose, pair lists by S block, work items) on 1, 3 and 8 host
    threads: identical tables (no GPU needed: viso_ba_debug_host_build)"""
    import viso_b200 as vb
    for kw in (dict(config_id=3, n_poses=40, n_points=9000, target_edges=60000), dict(config_id=4, n_poses=120, n_points=12000, target_edges=90000)):
        cid = kw.pop("config_id")
        p = vb.synth_config(cid, **kw)
        ba = vb.BundleAdjuster(len(p["points_init"]), len(p["poses_init"]), len(p["edge_pose"]), path=2)
        ba.set_problem(p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
        sums = {t: ba.host_build(t)[1] for t in (1, 3, 8)}
        assert len(set(sums.values())) == 1, sums
        ba.close()
