"""CPU test of the large reduced-system solver's plan (viso_b200/csrc/ba_solve_plan.h): tests/cpp/solve_plan_test.cpp runs
the plan through a sequential emulation of the kernel in ba_dense.cu (same 64-column steps, same row ranges, NaN outside
the zeroed envelope) and compares with the known solution of a block-banded SPD system.  Reference concept:
LinearSolverDense on the Schur-reduced pose matrix, reference include/bundle_adjuster.h:113-114."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "_build", "solve_plan_test")


@pytest.fixture(scope="module")
def exe():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "viso_b200", "csrc"),
                    os.path.join(ROOT, "tests", "cpp", "solve_plan_test.cpp"), "-o", EXE], check=True)
    return EXE


@pytest.mark.parametrize("nf,span,max_seg,force,segments", [
    (300, 20, 8, 0, 3),     # auto: interiors at least twice a separator
    (300, 20, 1, 0, 0),     # partition switched off: one envelope Cholesky
    (120, 10, 4, 1, 4),     # forced: interiors smaller than one 64-column block
    (40, 3, 8, 1, 2),       # forced on a tiny chain
    (500, 25, 8, 0, 5),     # the n = 2994 GPU parity test's shape
    (700, 33, 8, 0, 3),     # span just above 32 poses: separators of 64 poses
    (64, 63, 8, 1, 0),      # dense: nothing to cut
    (200, 199, 8, 0, 0),    # BASELINE configs[2]: dense S of 199 free poses
])
def test_plan_through_the_kernel_emulation(exe, nf, span, max_seg, force, segments):
    res = subprocess.run([exe, str(nf), str(span), str(max_seg), str(force), "148"], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    assert " segments %d " % segments in res.stdout, res.stdout


def test_plan_with_few_ctas(exe):
    # fewer CTAs than segments wanted (a small GPU): the plan caps the segment count
    res = subprocess.run([exe, "400", "8", "8", "1", "3"], capture_output=True, text=True)
    assert res.returncode == 0 and " segments 3 " in res.stdout, res.stdout + res.stderr
