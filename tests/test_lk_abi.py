"""CPU-only checks of the Lucas-Kanade boundary: libviso_ba.so exports every symbol include/viso_lk.h
declares, the Python mirror covers the whole header, and without a GPU the handle fails loudly."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "viso_lk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(viso_lk_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(vb):
    from viso_b200 import lk
    declared = _declared_symbols()
    assert len(declared) == 13
    lib = C.CDLL(vb.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), "libviso_ba.so does not export " + name
    assert sorted(lk.LK_SYMBOLS) == declared
    out = subprocess.run(["nm", "-D", "--defined-only", vb.LIB_PATH], capture_output=True, text=True).stdout
    assert set(declared) <= set(re.findall(r" T (viso_lk_[a-z0-9_]+)", out))


def test_default_config_is_the_reference_constants(vb):
    from viso_b200 import lk
    cfg = lk.default_config()
    assert cfg.struct_size == C.sizeof(lk.LkConfig)
    assert cfg.half_patch_size == 4                  # include/viso.h:25
    assert cfg.align_iterations == 100               # src/viso.cpp:849
    assert cfg.flow_iterations == 10                 # src/viso.cpp:266
    assert cfg.photometric_error_thresh == 14400.0   # include/viso.h:26


def test_without_a_gpu_the_handle_fails_loudly(vb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from viso_b200 import lk
    with pytest.raises(lk.LkError) as e:
        lk.LKAligner()
    assert e.value.code == 3  # ERR_CUDA
