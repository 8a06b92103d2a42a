"""ctypes binding of the CPU oracle of the direct pose estimator (oracle/direct_oracle.c).
TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs."""
import ctypes as C

import numpy as np

import oracle_lib

_D = C.POINTER(C.c_double)
_I32 = C.POINTER(C.c_int32)


class Image(C.Structure):
    _fields_ = [("data", C.POINTER(C.c_uint8)), ("cols", C.c_int32), ("rows", C.c_int32), ("step", C.c_int32)]


class Cfg(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("half_patch_size", C.c_int32), ("max_iterations", C.c_int32), ("delta_thresh", C.c_double),
                ("reference_termination", C.c_int32), ("reserved", C.c_int32)]


class LevelStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("n_good", C.c_int32), ("stop_reason", C.c_int32), ("reserved", C.c_int32),
                ("cost", C.c_double), ("last_cost", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        oracle_lib.build_oracle()
        L = C.CDLL(oracle_lib.ORACLE_LIB)
        L.direct_default_cfg.argtypes = [C.POINTER(Cfg)]
        L.direct_pixel_value.argtypes = [C.POINTER(Image), C.c_double, C.c_double]
        L.direct_pixel_value.restype = C.c_double
        L.direct_gradient.argtypes = [C.POINTER(Image), C.c_double, C.c_double, _D]
        L.direct_project.argtypes = [C.POINTER(Cfg), _D, _D, C.c_double, _D]
        L.direct_dpixel_dxi.argtypes = [C.POINTER(Cfg), _D, _D, _D, C.c_double, _D]
        L.direct_linearize.argtypes = [C.POINTER(Cfg), C.c_int32, C.POINTER(Image), C.POINTER(Image), _D, _D, _D, C.c_int32,
                                       _D, _D, _D, _I32]
        L.direct_single_layer.argtypes = [C.POINTER(Cfg), C.c_int32, C.POINTER(Image), C.POINTER(Image), _D, _D, _D, C.c_int32,
                                          C.POINTER(LevelStats)]
        L.direct_multi_layer.argtypes = [C.POINTER(Cfg), C.POINTER(Image), C.POINTER(Image), _D, _D, _D, C.c_int32,
                                         C.POINTER(LevelStats)]
        _lib = L
    return _lib


def cfg(**kw):
    c = Cfg()
    lib().direct_default_cfg(C.byref(c))
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def image(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    im = Image(a.ctypes.data_as(C.POINTER(C.c_uint8)), a.shape[1], a.shape[0], a.strides[0])
    im._keep = a
    return im


def _images(pyr):
    arr = (Image * len(pyr))()
    keep = []
    for i, a in enumerate(pyr):
        im = image(a)
        keep.append(im._keep)
        arr[i] = im
    arr._keep = keep
    return arr


def _d(a):
    return a.ctypes.data_as(_D)


def pixel_value(a, x, y):
    return lib().direct_pixel_value(C.byref(image(a)), x, y)


def gradient(a, u, v):
    g = np.zeros(2)
    lib().direct_gradient(C.byref(image(a)), u, v, _d(g))
    return g


def linearize(c, level, ref, cur, ref_pose, T21, points):
    pts = np.ascontiguousarray(points, dtype=np.float64)
    rp, tp = np.ascontiguousarray(ref_pose, dtype=np.float64), np.ascontiguousarray(T21, dtype=np.float64)
    H, b, cost, ng = np.zeros((6, 6)), np.zeros(6), C.c_double(0), C.c_int32(0)
    r, q = image(ref), image(cur)
    lib().direct_linearize(C.byref(c), level, C.byref(r), C.byref(q), _d(rp), _d(tp), _d(pts), len(pts), _d(H), _d(b),
                           C.byref(cost), C.byref(ng))
    return H, b, cost.value, ng.value


def single_layer(c, level, ref, cur, ref_pose, T21, points):
    pts = np.ascontiguousarray(points, dtype=np.float64)
    rp, tp = np.ascontiguousarray(ref_pose, dtype=np.float64), np.array(T21, dtype=np.float64)
    st = LevelStats()
    r, q = image(ref), image(cur)
    lib().direct_single_layer(C.byref(c), level, C.byref(r), C.byref(q), _d(rp), _d(tp), _d(pts), len(pts), C.byref(st))
    return tp, st


def multi_layer(c, ref_pyr, cur_pyr, ref_pose, T21, points):
    pts = np.ascontiguousarray(points, dtype=np.float64)
    rp, tp = np.ascontiguousarray(ref_pose, dtype=np.float64), np.array(T21, dtype=np.float64)
    st = (LevelStats * 4)()
    r, q = _images(ref_pyr), _images(cur_pyr)
    lib().direct_multi_layer(C.byref(c), r, q, _d(rp), _d(tp), _d(pts), len(pts), st)
    return tp, list(st)
