"""CPU-only checks of the direct pose estimator's boundary: libviso_ba.so exports every symbol
include/viso_direct.h declares, the Python mirror covers the whole header, and without a GPU the
handle fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "viso_direct.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(viso_direct_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(vb):
    from viso_b200 import direct
    declared = _declared_symbols()
    assert len(declared) == 12
    lib = C.CDLL(vb.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), "libviso_ba.so does not export " + name
    assert sorted(direct.DIRECT_SYMBOLS) == declared
    out = subprocess.run(["nm", "-D", "--defined-only", vb.LIB_PATH], capture_output=True, text=True).stdout
    assert set(declared) <= set(re.findall(r" T (viso_direct_[a-z0-9_]+)", out))


def test_default_config_is_the_reference_constants(vb):
    from viso_b200 import direct
    cfg = direct.default_config()
    assert cfg.struct_size == C.sizeof(direct.DirectConfig)
    assert (cfg.fx, cfg.fy, cfg.cx, cfg.cy) == (517.3, 516.5, 325.1, 249.7)  # src/main.cpp:14-17
    assert cfg.half_patch_size == 4          # include/viso.h:25
    assert cfg.max_iterations == 100         # src/viso.cpp:671
    assert cfg.delta_thresh == 0.005         # src/viso.cpp:668
    assert cfg.reference_termination == 1


def test_bad_config_is_rejected(vb):
    from viso_b200 import direct
    lib = direct._load()
    cfg = direct.default_config()
    cfg.struct_size = 8
    st = C.c_int32(0)
    assert not lib.viso_direct_create(C.byref(cfg), C.byref(st))
    assert st.value == 6  # ERR_ARG


def test_without_a_gpu_the_handle_fails_loudly(vb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from viso_b200 import direct
    with pytest.raises(direct.DirectError) as e:
        direct.DirectPoseEstimator()
    assert e.value.code == 3  # ERR_CUDA
