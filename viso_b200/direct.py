"""Python mirror of the direct pose estimator's C ABI (include/viso_direct.h) — SURVEY.md 8 row f3.

Mirrors Viso::DirectPoseEstimationSingleLayer / MultiLayer (reference src/viso.cpp:661-766): same
argument meaning (level, current frame's pyramid, T21 in/out), last_frame and the map points set on
the handle the way Viso holds them as members.  ctypes over libviso_ba.so; no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import api as _api

LEVELS = 4
_D = C.POINTER(C.c_double)
_U8 = C.POINTER(C.c_uint8)
_I32 = C.POINTER(C.c_int32)


class DirectConfig(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("device", C.c_int32), ("max_points", C.c_int32), ("max_frames", C.c_int32),
                ("max_width", C.c_int32), ("max_height", C.c_int32),
                ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("half_patch_size", C.c_int32), ("max_iterations", C.c_int32), ("delta_thresh", C.c_double),
                ("reference_termination", C.c_int32), ("reserved", C.c_int32)]


class LevelStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("n_good", C.c_int32), ("stop_reason", C.c_int32), ("reserved", C.c_int32),
                ("cost", C.c_double), ("last_cost", C.c_double)]


_P = C.c_void_p
DIRECT_SYMBOLS = {
    "viso_direct_default_config": (None, [C.POINTER(DirectConfig)]),
    "viso_direct_create": (_P, [C.POINTER(DirectConfig), _I32]),
    "viso_direct_destroy": (None, [_P]),
    "viso_direct_last_error": (C.c_char_p, [_P]),
    "viso_direct_kernel_launches": (C.c_int64, [_P]),
    "viso_direct_set_image": (C.c_int32, [_P, C.c_int32, C.c_int32, C.c_int32, _U8, C.c_int32, C.c_int32, C.c_int32]),
    "viso_direct_set_frame": (C.c_int32, [_P, C.c_int32, C.c_int32, _U8, C.c_int32, C.c_int32, C.c_int32]),
    "viso_direct_set_points": (C.c_int32, [_P, C.c_int32, _D, C.c_int32]),
    "viso_direct_set_reference_pose": (C.c_int32, [_P, C.c_int32, _D]),
    "viso_direct_estimate_level": (C.c_int32, [_P, C.c_int32, C.c_int32, _D, C.POINTER(LevelStats)]),
    "viso_direct_estimate": (C.c_int32, [_P, C.c_int32, _D, C.POINTER(LevelStats), C.POINTER(C.c_float)]),
    "viso_direct_linearize": (C.c_int32, [_P, C.c_int32, C.c_int32, _D, _D, _D, _D, _I32]),
}

_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = _api.load_library()
        for name, (res, args) in DIRECT_SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def default_config():
    cfg = DirectConfig()
    _load().viso_direct_default_config(C.byref(cfg))
    return cfg


class DirectError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("viso_direct status %d: %s" % (code, msg))
        self.code = code


class DirectPoseEstimator:
    """One handle = device images + map points + one stream.  `frames` > 1 batches independent
    (last_frame, current_frame, T21) problems into one launch."""

    def __init__(self, max_points=4096, frames=1, device=0, width=640, height=480, **overrides):
        lib = _load()
        cfg = default_config()
        cfg.max_points, cfg.max_frames, cfg.device, cfg.max_width, cfg.max_height = max_points, frames, device, width, height
        for k, v in overrides.items():
            if not hasattr(cfg, k):
                raise TypeError("unknown config field %r" % k)
            setattr(cfg, k, v)
        self.cfg = cfg
        st = C.c_int32(0)
        self._h = lib.viso_direct_create(C.byref(cfg), C.byref(st))
        if not self._h:
            raise DirectError(st.value, "viso_direct_create failed")
        if st.value != 0:
            msg = lib.viso_direct_last_error(self._h).decode()
            lib.viso_direct_destroy(self._h)
            self._h = None
            raise DirectError(st.value, msg)
        self.frames = frames

    def close(self):
        if getattr(self, "_h", None):
            _load().viso_direct_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, code):
        if code != 0:
            raise DirectError(code, _load().viso_direct_last_error(self._h).decode())

    def set_pyramid(self, which, pyramid, frame=0):
        """which: 0 = last_frame, 1 = current_frame; pyramid: list of 4 uint8 images (Keyframe::Pyramids())."""
        for level, im in enumerate(pyramid):
            im = np.ascontiguousarray(im, dtype=np.uint8)
            self._check(_load().viso_direct_set_image(self._h, frame, which, level, im.ctypes.data_as(_U8),
                                                      im.shape[1], im.shape[0], im.strides[0]))

    def set_frame(self, which, image, frame=0):
        """Keyframe(cv::Mat) (keyframe.h:29-45): level 0 = image, levels 1..3 by pyrDown on the device"""
        im = np.ascontiguousarray(image, dtype=np.uint8)
        self._check(_load().viso_direct_set_frame(self._h, frame, which, im.ctypes.data_as(_U8), im.shape[1], im.shape[0], im.strides[0]))

    def set_points(self, xyz, frame=-1):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        self._check(_load().viso_direct_set_points(self._h, frame, xyz.ctypes.data_as(_D), len(xyz)))

    def set_reference_pose(self, pose, frame=0):
        pose = np.ascontiguousarray(pose, dtype=np.float64).reshape(7)
        self._check(_load().viso_direct_set_reference_pose(self._h, frame, pose.ctypes.data_as(_D)))

    def DirectPoseEstimationSingleLayer(self, level, T21):
        """reference src/viso.cpp:661-758.  T21: [7] or [frames][7]; returns (refined, [LevelStats per frame])."""
        poses = np.array(T21, dtype=np.float64).reshape(-1, 7).copy()
        st = (LevelStats * len(poses))()
        self._check(_load().viso_direct_estimate_level(self._h, level, len(poses), poses.ctypes.data_as(_D), st))
        return poses.reshape(np.shape(T21)), list(st)

    def DirectPoseEstimationMultiLayer(self, T21):
        """reference src/viso.cpp:760-766.  Returns (refined, stats[frame][level], device_ms)."""
        poses = np.array(T21, dtype=np.float64).reshape(-1, 7).copy()
        st = (LevelStats * (LEVELS * len(poses)))()
        ms = C.c_float(0)
        self._check(_load().viso_direct_estimate(self._h, len(poses), poses.ctypes.data_as(_D), st, C.byref(ms)))
        stats = [[st[LEVELS * f + l] for l in range(LEVELS)] for f in range(len(poses))]
        return poses.reshape(np.shape(T21)), stats, float(ms.value)

    def linearize(self, level, T21):
        """H [frames][6][6], b [frames][6], cost [frames], n_good [frames] of one linearisation (viso.cpp:688-733)."""
        poses = np.array(T21, dtype=np.float64).reshape(-1, 7).copy()
        n = len(poses)
        H, b, cost = np.zeros((n, 6, 6)), np.zeros((n, 6)), np.zeros(n)
        ng = np.zeros(n, dtype=np.int32)
        self._check(_load().viso_direct_linearize(self._h, level, n, poses.ctypes.data_as(_D), H.ctypes.data_as(_D),
                                                  b.ctypes.data_as(_D), cost.ctypes.data_as(_D), ng.ctypes.data_as(_I32)))
        return H, b, cost, ng

    def kernel_launches(self):
        return int(_load().viso_direct_kernel_launches(self._h))
