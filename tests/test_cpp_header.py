"""Source-compat test (last row): a C++ caller that uses the reference's class
surface — BundleAdjuster(), AddPoint, AddPose, AddEdge (reference include/bundle_adjuster.h:111,
:169-173) — builds against include/bundle_adjuster.h with plain g++ (no Eigen / Sophus / g2o)
and links against libviso_ba.so."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "drop_in.cpp")
OUT = os.path.join(ROOT, "tests", "cpp", "_build")
EXE = os.path.join(OUT, "drop_in")


def _build(vb):
    os.makedirs(OUT, exist_ok=True)
    libdir = os.path.dirname(vb.LIB_PATH)
    cmd = ["g++", "-std=c++11", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), SRC, "-o", EXE,
           "-L", libdir, "-l:libviso_ba.so", "-Wl,-rpath," + libdir]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return EXE


def test_reference_surface_compiles_and_stages_without_gpu(vb):
    exe = _build(vb)
    res = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, (res.returncode, res.stdout, res.stderr)
    assert "staging ok" in res.stdout


@pytest.mark.gpu
def test_cpp_caller_optimizes_on_the_gpu(vb):
    exe = _build(vb)
    res = subprocess.run([exe, "--optimize"], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, (res.returncode, res.stdout, res.stderr)
