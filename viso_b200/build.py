"""Build libviso_ba.so (CUDA kernels + C ABI host) in-tree with nvcc for sm_100a, and the host-only
synthetic-problem generator libviso_synth.so (g++, no CUDA) beside it.

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.  nvcc
cross-compiles without a GPU, so this runs in the CPU container too.
"""
import hashlib
import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libviso_ba.so")
STAMP = os.path.join(HERE, ".libviso_ba.stamp")
SYNTH_LIB = os.path.join(HERE, "libviso_synth.so")
SOURCES = ["ba_kernels.cu", "ba_window.cu", "ba_dense.cu", "ba_setup.cu", "ba_photo.cu", "ba_static_cost.cu", "direct_pose.cu", "lk_align.cu", "pyramid.cu", "viso_ba.cpp"]
SYNTH_SOURCES = ["synth.cpp"]
HEADERS = sorted(glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh"))) + [os.path.join(ROOT, "include", h) for h in ("viso_ba.h", "viso_ba_synth.h", "viso_direct.h", "viso_lk.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O2,-Wall", "-shared", "-cudart", "static",
]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: the CUDA backend cannot be built (there is no CPU fallback)")


def _digest():
    h = hashlib.sha256()
    for f in [os.path.join(CSRC, s) for s in SOURCES + SYNTH_SOURCES] + HEADERS + [os.path.abspath(__file__)]:
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def build(force=False, verbose=False):
    """Compile every CUDA source of the package; returns the path of the shared library."""
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(SYNTH_LIB) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == dig:
                return LIB
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
    cmd += [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB, "-ldl"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose:
        sys.stderr.write(res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    gxx = shutil.which("g++") or "g++"
    res = subprocess.run([gxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall"] + [os.path.join(CSRC, s) for s in SYNTH_SOURCES]
                         + ["-o", SYNTH_LIB], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("g++ failed (libviso_synth.so):\n" + res.stdout + res.stderr)
    with open(STAMP, "w") as fh:
        fh.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
