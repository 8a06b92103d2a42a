"""ctypes binding of the CPU oracle (oracle/ba_oracle.c).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "_build", "libba_oracle.so")


class OracleCfg(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("huber_delta", C.c_double), ("tau", C.c_double), ("lambda_init", C.c_double),
                ("max_trials", C.c_int32), ("threads", C.c_int32)]


class OracleIterStats(C.Structure):
    _fields_ = [("cost_before", C.c_double), ("cost_after", C.c_double), ("lambda_", C.c_double),
                ("trials", C.c_int32), ("accepted", C.c_int32)]


_D = C.POINTER(C.c_double)
_I32 = C.POINTER(C.c_int32)
_U8 = C.POINTER(C.c_uint8)
_F = C.POINTER(C.c_float)
_lib = None


def build_oracle(force=False):
    src = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".c", ".h")) or f == "Makefile"]
    if (not force and os.path.exists(ORACLE_LIB)
            and all(os.path.getmtime(ORACLE_LIB) >= os.path.getmtime(s) for s in src)):
        return ORACLE_LIB
    subprocess.run(["make", "-C", ORACLE_DIR, "-B"], check=True, capture_output=True)
    return ORACLE_LIB


def lib():
    global _lib
    if _lib is None:
        build_oracle()
        L = C.CDLL(ORACLE_LIB)
        L.oracle_default_cfg.argtypes = [C.POINTER(OracleCfg)]
        L.oracle_quat_to_R.argtypes = [_D, _D]
        L.oracle_project.argtypes = [C.POINTER(OracleCfg), _D, _D, _D, _D]
        L.oracle_jac_pose.argtypes = [C.POINTER(OracleCfg), _D, _D]
        L.oracle_jac_point.argtypes = [C.POINTER(OracleCfg), _D, _D, _D]
        L.oracle_se3_exp.argtypes = [_D, _D]
        L.oracle_pose_oplus.argtypes = [_D, _D]
        L.oracle_huber.argtypes = [C.c_double, C.c_double, _D]
        L.oracle_huber.restype = C.c_double
        L.oracle_cost.argtypes = [C.POINTER(OracleCfg), C.c_int32, _D, C.c_int32, _D, C.c_int64, _I32, _I32, _F]
        L.oracle_cost.restype = C.c_double
        L.oracle_reduced_system.argtypes = [C.POINTER(OracleCfg), C.c_int32, _D, _U8, C.c_int32, _D, C.c_int64,
                                            _I32, _I32, _F, C.c_double, _D, _D, _D, _I32]
        L.oracle_reduced_system.restype = C.c_int32
        L.oracle_optimize.argtypes = [C.POINTER(OracleCfg), C.c_int32, _D, _U8, C.c_int32, _D, C.c_int64,
                                      _I32, _I32, _F, C.c_int32, C.POINTER(OracleIterStats), _D]
        L.oracle_optimize.restype = C.c_int32
        L.oracle_cholesky_solve.argtypes = [C.c_int32, _D, _D, C.c_int32]
        L.oracle_cholesky_solve.restype = C.c_int32
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_D)


def cfg(K=(517.3, 516.5, 325.1, 249.7), huber_delta=0.0, tau=1e-5, lambda_init=0.0, max_trials=10, threads=1):
    c = OracleCfg()
    lib().oracle_default_cfg(C.byref(c))
    c.fx, c.fy, c.cx, c.cy = K
    c.huber_delta, c.tau, c.lambda_init, c.max_trials, c.threads = huber_delta, tau, lambda_init, max_trials, threads
    return c


def _prep(poses, points, edge_pose, edge_point, edge_uv, fixed):
    poses = np.array(poses, dtype=np.float64, order="C").reshape(-1, 7)
    points = np.array(points, dtype=np.float64, order="C").reshape(-1, 3)
    ep = np.ascontiguousarray(edge_pose, np.int32)
    el = np.ascontiguousarray(edge_point, np.int32)
    uv = np.ascontiguousarray(edge_uv, np.float32)
    fx = np.zeros(len(poses), np.uint8)
    if fixed is None:
        fx[0] = 1
    else:
        fx[:] = np.asarray(fixed, np.uint8)
    return poses, points, ep, el, uv, fx


def cost(c, poses, points, edge_pose, edge_point, edge_uv):
    poses, points, ep, el, uv, _ = _prep(poses, points, edge_pose, edge_point, edge_uv, None)
    return lib().oracle_cost(C.byref(c), len(poses), _d(poses), len(points), _d(points), len(ep),
                             ep.ctypes.data_as(_I32), el.ctypes.data_as(_I32), uv.ctypes.data_as(_F))


def reduced_system(c, poses, points, edge_pose, edge_point, edge_uv, lam, fixed=None):
    poses, points, ep, el, uv, fx = _prep(poses, points, edge_pose, edge_point, edge_uv, fixed)
    n = 6 * int((fx == 0).sum())
    S, bs = np.zeros((n, n)), np.zeros(n)
    md = C.c_double()
    fi = np.zeros(len(poses), np.int32)
    r = lib().oracle_reduced_system(C.byref(c), len(poses), _d(poses), fx.ctypes.data_as(_U8), len(points), _d(points),
                                    len(ep), ep.ctypes.data_as(_I32), el.ctypes.data_as(_I32), uv.ctypes.data_as(_F),
                                    lam, _d(S), _d(bs), C.byref(md), fi.ctypes.data_as(_I32))
    assert r == n, r
    return S, bs, md.value


def optimize(c, poses, points, edge_pose, edge_point, edge_uv, iterations, fixed=None):
    """Returns (poses, points, final_cost, iterations_run, [stats])."""
    poses, points, ep, el, uv, fx = _prep(poses, points, edge_pose, edge_point, edge_uv, fixed)
    stats = (OracleIterStats * max(1, iterations))()
    fc = C.c_double()
    r = lib().oracle_optimize(C.byref(c), len(poses), _d(poses), fx.ctypes.data_as(_U8), len(points), _d(points),
                              len(ep), ep.ctypes.data_as(_I32), el.ctypes.data_as(_I32), uv.ctypes.data_as(_F),
                              iterations, stats, C.byref(fc))
    assert r >= 0, r
    return poses, points, fc.value, r, list(stats)[:r]
