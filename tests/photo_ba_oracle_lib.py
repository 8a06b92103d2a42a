"""ctypes binding of the CPU oracle of photometric bundle adjustment (oracle/photo_ba_oracle.c).
TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs."""
import ctypes as C

import numpy as np

import direct_oracle_lib as dorc
import oracle_lib

_D = C.POINTER(C.c_double)
_I32 = C.POINTER(C.c_int32)
_U8 = C.POINTER(C.c_uint8)
_IM = C.POINTER(dorc.Image)
_CFG = C.POINTER(oracle_lib.OracleCfg)
_lib = None


def lib():
    global _lib
    if _lib is None:
        oracle_lib.build_oracle()
        L = C.CDLL(oracle_lib.ORACLE_LIB)
        L.photo_ba_cost.argtypes = [_CFG, _IM, C.c_int32, _D, _D, _D, C.c_int64, _I32, _I32, _I32]
        L.photo_ba_cost.restype = C.c_double
        L.photo_ba_edge_jacobian.argtypes = [_CFG, _IM, _IM, _D, _D, _D, _D, _D, _D]
        L.photo_ba_edge_jacobian.restype = C.c_int32
        L.photo_ba_optimize.argtypes = [_CFG, _IM, C.c_int32, _D, _U8, C.c_int32, _D, C.c_int64, _I32, _I32, _I32, C.c_int32,
                                        C.POINTER(oracle_lib.OracleIterStats), _D]
        L.photo_ba_optimize.restype = C.c_int32
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_D)


def _i(a):
    return a.ctypes.data_as(_I32)


def cost(c, images, poses, src_poses, points, e_point, e_src, e_dst):
    poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
    src_poses = np.ascontiguousarray(src_poses, np.float64).reshape(-1, 7)
    points = np.ascontiguousarray(points, np.float64).reshape(-1, 3)
    el, es, ed = (np.ascontiguousarray(a, np.int32) for a in (e_point, e_src, e_dst))
    ims = dorc._images(images)
    return lib().photo_ba_cost(C.byref(c), ims, len(poses), _d(poses), _d(src_poses), _d(points), len(el), _i(el), _i(es), _i(ed))


def edge_jacobian(c, src_img, tgt_img, src_pose, dst_pose, p):
    """-> (valid, r[16], Jp[16][6], Jl[16][3])"""
    r, Jp, Jl = np.zeros(16), np.zeros((16, 6)), np.zeros((16, 3))
    sp, dp, pp = (np.ascontiguousarray(a, np.float64) for a in (src_pose, dst_pose, p))
    si, ti = dorc.image(src_img), dorc.image(tgt_img)
    ok = lib().photo_ba_edge_jacobian(C.byref(c), C.byref(si), C.byref(ti), _d(sp), _d(dp), _d(pp), _d(r), _d(Jp), _d(Jl))
    return bool(ok), r, Jp, Jl


def optimize(c, images, poses, points, e_point, e_src, e_dst, iterations, fixed=None):
    """-> (poses, points, final_cost, iterations_run, [stats]); the source poses are frozen at the incoming poses"""
    poses = np.array(poses, dtype=np.float64, order="C").reshape(-1, 7)
    points = np.array(points, dtype=np.float64, order="C").reshape(-1, 3)
    el, es, ed = (np.ascontiguousarray(a, np.int32) for a in (e_point, e_src, e_dst))
    fx = np.zeros(len(poses), np.uint8)
    if fixed is None:
        fx[0] = 1
    else:
        fx[:] = np.asarray(fixed, np.uint8)
    ims = dorc._images(images)
    stats = (oracle_lib.OracleIterStats * max(1, iterations))()
    fc = C.c_double()
    r = lib().photo_ba_optimize(C.byref(c), ims, len(poses), _d(poses), fx.ctypes.data_as(_U8), len(points), _d(points), len(el),
                                _i(el), _i(es), _i(ed), iterations, stats, C.byref(fc))
    assert r >= 0, r
    return poses, points, fc.value, r, list(stats)[:r]
