"""SURVEY section 8 row f1 — the C++ wiring into viso's VO loop, integration/viso_ba_glue.h: Map -> (point, pose, pixel)
extraction, the RingBuffer-shaped window of the last N keyframes, Optimize, write-back (reference call sites
src/viso.cpp:77-98 and :113-138; containers include/map.h:15-19, include/map_point.h:21-31, include/ring_buffer.h:10-42).
OpenCV / Eigen / Sophus are absent here, so the header is compiled against plain structs that carry the reference's
accessor names (tests/cpp/glue_test.cpp) with plain g++ -std=c++11, like the reference (CMakeLists.txt:3)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "glue_test.cpp")
OUT = os.path.join(ROOT, "tests", "cpp", "_build")
EXE = os.path.join(OUT, "glue_test")


def _build(vb):
    os.makedirs(OUT, exist_ok=True)
    libdir = os.path.dirname(vb.LIB_PATH)
    cmd = ["g++", "-std=c++11", "-O2", "-Wall", "-Werror", SRC, "-o", EXE, "-L", libdir, "-l:libviso_ba.so", "-Wl,-rpath," + libdir]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return EXE


def test_glue_compiles_and_extracts_the_window_without_gpu(vb):
    res = subprocess.run([_build(vb), "extract"], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, (res.returncode, res.stdout, res.stderr)
    assert "extract ok: 4 keyframes in the window" in res.stdout


@pytest.mark.gpu
def test_glue_adjusts_the_window_on_the_gpu_and_writes_back(vb):
    res = subprocess.run([_build(vb), "optimize"], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, (res.returncode, res.stdout, res.stderr)
    assert "optimize ok" in res.stdout
