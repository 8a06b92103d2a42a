"""ctypes binding of the CPU oracle of the Lucas-Kanade alignment (oracle/lk_oracle.c).
TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs."""
import ctypes as C

import numpy as np

import direct_oracle_lib as dorc
import oracle_lib

_D = C.POINTER(C.c_double)
_F = C.POINTER(C.c_float)
_U8 = C.POINTER(C.c_uint8)
_I32 = C.POINTER(C.c_int32)
_IM = C.POINTER(dorc.Image)


class Cfg(C.Structure):
    _fields_ = [("half_patch_size", C.c_int32), ("align_iterations", C.c_int32), ("flow_iterations", C.c_int32),
                ("reserved", C.c_int32), ("photometric_error_thresh", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        oracle_lib.build_oracle()
        L = C.CDLL(oracle_lib.ORACLE_LIB)
        L.lk_default_cfg.argtypes = [C.POINTER(Cfg)]
        L.lk_align_single.argtypes = [C.POINTER(Cfg), _IM, _I32, _IM, C.c_int32, _D, _D, _U8, C.c_int32]
        L.lk_align.argtypes = [C.POINTER(Cfg), _IM, _I32, _IM, C.c_int32, _D, _D, _U8]
        L.lk_flow_single.argtypes = [C.POINTER(Cfg), _IM, _IM, C.c_int32, _F, _F, C.c_int32, C.c_int32, _U8]
        L.lk_flow_multi.argtypes = [C.POINTER(Cfg), _IM, _IM, C.c_int32, _F, _F, C.c_int32, C.c_int32, _U8]
        L.photo_edge_errors.argtypes = [_IM, C.c_int32, _D, _I32, _I32, _D, _D, _D, _D, _U8]
        _lib = L
    return _lib


def cfg(**kw):
    c = Cfg()
    lib().lk_default_cfg(C.byref(c))
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def _slots(pyramids):
    """[n_slots] pyramids (4 levels each) -> flat Image array [n_slots * 4]"""
    flat = [im for pyr in pyramids for im in pyr]
    return dorc._images(flat)


def align(c, ref_pyramids, ref_slot, cur_pyr, uv_ref, uv_cur, level=None):
    uv_ref = np.ascontiguousarray(uv_ref, dtype=np.float64).reshape(-1, 2)
    uv_cur = np.array(uv_cur, dtype=np.float64).reshape(-1, 2).copy()
    n = len(uv_ref)
    rs = np.ascontiguousarray(np.broadcast_to(np.asarray(ref_slot, dtype=np.int32), (n,)))
    ok = np.zeros(n, dtype=np.uint8)
    refs, cur = _slots(ref_pyramids), dorc._images(cur_pyr)
    args = (C.byref(c), refs, rs.ctypes.data_as(_I32), cur, n, uv_ref.ctypes.data_as(_D), uv_cur.ctypes.data_as(_D), ok.ctypes.data_as(_U8))
    if level is None:
        lib().lk_align(*args)
    else:
        lib().lk_align_single(*args, level)
    return uv_cur, ok.astype(bool)


def flow(c, pyr1, pyr2, kp1, kp2=None, inverse=False, level=None):
    kp1 = np.ascontiguousarray(kp1, dtype=np.float32).reshape(-1, 2)
    have = kp2 is not None and len(kp2) > 0
    kp2 = np.array(kp2, dtype=np.float32).reshape(-1, 2).copy() if have else np.zeros_like(kp1)
    ok = np.zeros(len(kp1), dtype=np.uint8)
    if level is None:
        a, b = dorc._images(pyr1), dorc._images(pyr2)
        lib().lk_flow_multi(C.byref(c), a, b, len(kp1), kp1.ctypes.data_as(_F), kp2.ctypes.data_as(_F), int(have), int(inverse), ok.ctypes.data_as(_U8))
    else:
        a, b = dorc.image(pyr1[level]), dorc.image(pyr2[level])
        lib().lk_flow_single(C.byref(c), C.byref(a), C.byref(b), len(kp1), kp1.ctypes.data_as(_F), kp2.ctypes.data_as(_F), int(have), int(inverse),
                             ok.ctypes.data_as(_U8))
    return kp2, ok.astype(bool)


def photo_edge_errors(pyramids, points, src_slot, tgt_slot, frame_poses, vertex_poses, K=(517.3, 516.5, 325.1, 249.7)):
    """include/bundle_adjuster.h:70-101 for n edges"""
    pts = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
    n = len(pts)
    ss = np.ascontiguousarray(np.broadcast_to(np.asarray(src_slot, dtype=np.int32), (n,)))
    ts = np.ascontiguousarray(np.broadcast_to(np.asarray(tgt_slot, dtype=np.int32), (n,)))
    fp = np.ascontiguousarray(frame_poses, dtype=np.float64)
    vp = np.ascontiguousarray(vertex_poses, dtype=np.float64)
    Kc = np.array(K, dtype=np.float64)
    err, lvl = np.zeros((n, 16)), np.zeros(n, dtype=np.uint8)
    imgs = _slots(pyramids)
    lib().photo_edge_errors(imgs, n, pts.ctypes.data_as(_D), ss.ctypes.data_as(_I32), ts.ctypes.data_as(_I32), fp.ctypes.data_as(_D),
                            vp.ctypes.data_as(_D), Kc.ctypes.data_as(_D), err.ctypes.data_as(_D), lvl.ctypes.data_as(_U8))
    return err, lvl
