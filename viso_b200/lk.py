"""Python mirror of the Lucas-Kanade patch alignment's C ABI (include/viso_lk.h) — SURVEY.md 8 row f4.

Mirrors Viso::LKAlignmentSingle / the level loop of Viso::LKAlignment (reference src/viso.cpp:845-925,
:828-830) and Viso::OpticalFlowSingleLevel / MultiLevel (src/viso.cpp:259-391): same argument meaning,
frames referred to by image slot instead of Keyframe::Ptr.  ctypes over libviso_ba.so; no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import api as _api

LEVELS = 4
_D = C.POINTER(C.c_double)
_F = C.POINTER(C.c_float)
_U8 = C.POINTER(C.c_uint8)
_I32 = C.POINTER(C.c_int32)
_P = C.c_void_p


class LkConfig(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("device", C.c_int32), ("max_keypoints", C.c_int32), ("max_slots", C.c_int32),
                ("max_width", C.c_int32), ("max_height", C.c_int32), ("half_patch_size", C.c_int32),
                ("align_iterations", C.c_int32), ("flow_iterations", C.c_int32), ("reserved", C.c_int32),
                ("photometric_error_thresh", C.c_double)]


LK_SYMBOLS = {
    "viso_lk_default_config": (None, [C.POINTER(LkConfig)]),
    "viso_lk_create": (_P, [C.POINTER(LkConfig), _I32]),
    "viso_lk_destroy": (None, [_P]),
    "viso_lk_last_error": (C.c_char_p, [_P]),
    "viso_lk_kernel_launches": (C.c_int64, [_P]),
    "viso_lk_set_image": (C.c_int32, [_P, C.c_int32, C.c_int32, _U8, C.c_int32, C.c_int32, C.c_int32]),
    "viso_lk_set_frame": (C.c_int32, [_P, C.c_int32, _U8, C.c_int32, C.c_int32, C.c_int32]),
    "viso_lk_get_image": (C.c_int32, [_P, C.c_int32, C.c_int32, _U8, C.c_int64, _I32, _I32]),
    "viso_lk_align": (C.c_int32, [_P, C.c_int32, _I32, C.c_int32, _D, _D, _U8, _F]),
    "viso_lk_align_level": (C.c_int32, [_P, C.c_int32, C.c_int32, _I32, C.c_int32, _D, _D, _U8]),
    "viso_lk_optical_flow": (C.c_int32, [_P, C.c_int32, C.c_int32, C.c_int32, _F, _F, C.c_int32, C.c_int32, _U8, _F]),
    "viso_lk_edge_direct_projection": (C.c_int32, [_P, C.c_int32, _D, _I32, _I32, _D, _D, C.c_double, C.c_double, C.c_double, C.c_double,
                                                   _D, _U8, _F]),
    "viso_lk_optical_flow_level": (C.c_int32, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _F, _F, C.c_int32, C.c_int32, _U8]),
}

_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = _api.load_library()
        for name, (res, args) in LK_SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def default_config():
    cfg = LkConfig()
    _load().viso_lk_default_config(C.byref(cfg))
    return cfg


class LkError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("viso_lk status %d: %s" % (code, msg))
        self.code = code


class LKAligner:
    """One handle = device pyramids of the image slots + keypoint buffers + one stream."""

    def __init__(self, max_keypoints=4096, slots=8, device=0, width=640, height=480, **overrides):
        lib = _load()
        cfg = default_config()
        cfg.max_keypoints, cfg.max_slots, cfg.device, cfg.max_width, cfg.max_height = max_keypoints, slots, device, width, height
        for k, v in overrides.items():
            if not hasattr(cfg, k):
                raise TypeError("unknown config field %r" % k)
            setattr(cfg, k, v)
        self.cfg = cfg
        st = C.c_int32(0)
        self._h = lib.viso_lk_create(C.byref(cfg), C.byref(st))
        if not self._h:
            raise LkError(st.value, "viso_lk_create failed")
        if st.value != 0:
            msg = lib.viso_lk_last_error(self._h).decode()
            lib.viso_lk_destroy(self._h)
            self._h = None
            raise LkError(st.value, msg)

    def close(self):
        if getattr(self, "_h", None):
            _load().viso_lk_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, code):
        if code != 0:
            raise LkError(code, _load().viso_lk_last_error(self._h).decode())

    def set_pyramid(self, slot, pyramid):
        for level, im in enumerate(pyramid):
            im = np.ascontiguousarray(im, dtype=np.uint8)
            self._check(_load().viso_lk_set_image(self._h, slot, level, im.ctypes.data_as(_U8), im.shape[1], im.shape[0], im.strides[0]))

    def set_frame(self, slot, image):
        """Keyframe(cv::Mat) (keyframe.h:29-45): level 0 = image, levels 1..3 by pyrDown on the device"""
        im = np.ascontiguousarray(image, dtype=np.uint8)
        self._check(_load().viso_lk_set_frame(self._h, slot, im.ctypes.data_as(_U8), im.shape[1], im.shape[0], im.strides[0]))

    def get_image(self, slot, level):
        """Keyframe::Pyramids()[level] of the slot, read back from the device"""
        out = np.zeros(self.cfg.max_width * self.cfg.max_height, dtype=np.uint8)
        cols, rows = C.c_int32(0), C.c_int32(0)
        self._check(_load().viso_lk_get_image(self._h, slot, level, out.ctypes.data_as(_U8), out.size, C.byref(cols), C.byref(rows)))
        return out[:cols.value * rows.value].reshape(rows.value, cols.value).copy()

    def _pairs(self, ref_slot, uv_ref, uv_cur):
        uv_ref = np.ascontiguousarray(uv_ref, dtype=np.float64).reshape(-1, 2)
        uv_cur = np.array(uv_cur, dtype=np.float64).reshape(-1, 2).copy()
        rs = np.ascontiguousarray(np.broadcast_to(np.asarray(ref_slot, dtype=np.int32), (len(uv_ref),)))
        return rs, uv_ref, uv_cur, np.zeros(len(uv_ref), dtype=np.uint8)

    def LKAlignment(self, ref_slot, cur_slot, uv_ref, uv_cur):
        """levels 3..0 of LKAlignmentSingle (viso.cpp:828-830).  Returns (uv_cur aligned [n][2], success [n] bool, device_ms);
        the reference's kp_before / kp_after are uv_cur_in[success] / uv_cur_out[success] (viso.cpp:822-843, :919-923)."""
        rs, uv_ref, uv_cur, ok = self._pairs(ref_slot, uv_ref, uv_cur)
        ms = C.c_float(0)
        self._check(_load().viso_lk_align(self._h, len(uv_ref), rs.ctypes.data_as(_I32), cur_slot, uv_ref.ctypes.data_as(_D),
                                          uv_cur.ctypes.data_as(_D), ok.ctypes.data_as(_U8), C.byref(ms)))
        return uv_cur, ok.astype(bool), float(ms.value)

    def LKAlignmentSingle(self, level, ref_slot, cur_slot, uv_ref, uv_cur):
        """viso.cpp:845-925 on one level"""
        rs, uv_ref, uv_cur, ok = self._pairs(ref_slot, uv_ref, uv_cur)
        self._check(_load().viso_lk_align_level(self._h, level, len(uv_ref), rs.ctypes.data_as(_I32), cur_slot,
                                                uv_ref.ctypes.data_as(_D), uv_cur.ctypes.data_as(_D), ok.ctypes.data_as(_U8)))
        return uv_cur, ok.astype(bool)

    def _kps(self, kp1, kp2):
        kp1 = np.ascontiguousarray(kp1, dtype=np.float32).reshape(-1, 2)
        have = kp2 is not None and len(kp2) > 0
        kp2 = np.array(kp2, dtype=np.float32).reshape(-1, 2).copy() if have else np.zeros_like(kp1)
        return kp1, kp2, int(have), np.zeros(len(kp1), dtype=np.uint8)

    def OpticalFlowMultiLevel(self, slot1, slot2, kp1, kp2=None, inverse=False):
        """viso.cpp:353-391.  kp2 None or empty = no initial guess.  Returns (kp2 [n][2] float32, success, device_ms)."""
        kp1, kp2, have, ok = self._kps(kp1, kp2)
        ms = C.c_float(0)
        self._check(_load().viso_lk_optical_flow(self._h, slot1, slot2, len(kp1), kp1.ctypes.data_as(_F), kp2.ctypes.data_as(_F),
                                                 have, int(inverse), ok.ctypes.data_as(_U8), C.byref(ms)))
        return kp2, ok.astype(bool), float(ms.value)

    def OpticalFlowSingleLevel(self, level, slot1, slot2, kp1, kp2=None, inverse=False):
        """viso.cpp:259-351 on pyramid level `level` of the two slots"""
        kp1, kp2, have, ok = self._kps(kp1, kp2)
        self._check(_load().viso_lk_optical_flow_level(self._h, level, slot1, slot2, len(kp1), kp1.ctypes.data_as(_F),
                                                       kp2.ctypes.data_as(_F), have, int(inverse), ok.ctypes.data_as(_U8)))
        return kp2, ok.astype(bool)

    def EdgeDirectProjection_computeError(self, points, src_slot, tgt_slot, frame_poses, vertex_poses, K=(517.3, 516.5, 325.1, 249.7)):
        """reference include/bundle_adjuster.h:70-101 for n edges.  Returns (errors [n][16], level [n] uint8, device_ms)."""
        pts = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
        n = len(pts)
        ss = np.ascontiguousarray(np.broadcast_to(np.asarray(src_slot, dtype=np.int32), (n,)))
        ts = np.ascontiguousarray(np.broadcast_to(np.asarray(tgt_slot, dtype=np.int32), (n,)))
        fp = np.zeros((self.cfg.max_slots, 7))
        vp = np.zeros((self.cfg.max_slots, 7))
        fp[:, 3] = vp[:, 3] = 1.0
        fp[:len(frame_poses)] = frame_poses
        vp[:len(vertex_poses)] = vertex_poses
        err, lvl, ms = np.zeros((n, 16)), np.zeros(n, dtype=np.uint8), C.c_float(0)
        self._check(_load().viso_lk_edge_direct_projection(self._h, n, pts.ctypes.data_as(_D), ss.ctypes.data_as(_I32), ts.ctypes.data_as(_I32),
                                                           fp.ctypes.data_as(_D), vp.ctypes.data_as(_D), K[0], K[1], K[2], K[3],
                                                           err.ctypes.data_as(_D), lvl.ctypes.data_as(_U8), C.byref(ms)))
        return err, lvl, float(ms.value)

    def kernel_launches(self):
        return int(_load().viso_lk_kernel_launches(self._h))
