"""ctypes binding of the C ABI in include/viso_ba.h and a Python mirror of the reference's
`BundleAdjuster` class (reference include/bundle_adjuster.h:108-187).

The names AddPoint / AddPose / AddEdge are the reference's (bundle_adjuster.h:169-173); the
reference bodies are empty and there is no Optimize or read-back, so the mirror
completes them the same way the C++ drop-in header include/bundle_adjuster.h does.

There is no CPU fallback: if libviso_ba.so is missing, or no sm_100 device is usable,
every compute call raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libviso_ba.so")
SYNTH_LIB_PATH = os.path.join(_HERE, "libviso_synth.so")  # host-only generator (include/viso_ba_synth.h)

OK = 0
STATUS_NAMES = {
    0: "OK", 1: "ERR_CAPACITY", 2: "ERR_INDEX", 3: "ERR_CUDA", 4: "ERR_NCCL",
    5: "ERR_NONFINITE", 6: "ERR_ARG", 7: "ERR_STATE", 8: "ERR_UNSUPPORTED",
}


class VisoBaError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("viso_ba status %d (%s): %s" % (status, STATUS_NAMES.get(status, "?"), message))
        self.status = status


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("device", C.c_int32),
        ("max_points", C.c_int64), ("max_poses", C.c_int64), ("max_edges", C.c_int64),
        ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
        ("huber_delta", C.c_double), ("lm_tau", C.c_double), ("lm_lambda_init", C.c_double),
        ("lm_max_trials", C.c_int32), ("verbose", C.c_int32),
        ("fix_first_pose", C.c_int32), ("use_cuda_graph", C.c_int32),
        ("path", C.c_int32), ("solver_segments", C.c_int32), ("host_structure_build", C.c_int32),
    ]


class IterStats(C.Structure):
    _fields_ = [("cost_before", C.c_double), ("cost_after", C.c_double), ("lambda_", C.c_double),
                ("trials", C.c_int32), ("accepted", C.c_int32)]


class Summary(C.Structure):
    _fields_ = [("initial_cost", C.c_double), ("final_cost", C.c_double), ("final_lambda", C.c_double),
                ("iterations", C.c_int32), ("trials", C.c_int32), ("terminated", C.c_int32),
                ("reserved", C.c_int32), ("device_ms", C.c_float), ("setup_ms", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class KernelTime(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("launches", C.c_int64), ("device_ms", C.c_double)]


class SynthSpec(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("layout", C.c_int32),
        ("n_poses", C.c_int64), ("n_points", C.c_int64), ("target_edges", C.c_int64),
        ("seed", C.c_uint64),
        ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
        ("width", C.c_double), ("height", C.c_double), ("margin", C.c_double),
        ("baseline", C.c_double), ("noise_px", C.c_double), ("outlier_frac", C.c_double),
        ("pose_rot_sigma", C.c_double), ("pose_trans_sigma", C.c_double), ("point_sigma", C.c_double),
        ("shuffle_edges", C.c_int32), ("reserved", C.c_int32),
    ]


# every symbol include/viso_ba.h declares: (restype, argtypes)
_P = C.c_void_p
_D = C.POINTER(C.c_double)
_I32 = C.POINTER(C.c_int32)
_I64 = C.POINTER(C.c_int64)
_F = C.POINTER(C.c_float)
SYMBOLS = {
    "viso_ba_version": (C.c_int32, []),
    "viso_ba_status_string": (C.c_char_p, [C.c_int32]),
    "viso_ba_default_config": (None, [C.POINTER(Config)]),
    "viso_ba_create": (C.c_int32, [C.POINTER(Config), C.POINTER(_P)]),
    "viso_ba_destroy": (None, [_P]),
    "viso_ba_last_error": (C.c_char_p, [_P]),
    "viso_ba_clear": (C.c_int32, [_P]),
    "viso_ba_add_point": (C.c_int64, [_P, _D]),
    "viso_ba_add_pose": (C.c_int64, [_P, _D]),
    "viso_ba_add_edge": (C.c_int32, [_P, C.c_int64, C.c_int64, C.c_float, C.c_float]),
    "viso_ba_set_problem": (C.c_int32, [_P, C.c_int64, _D, C.c_int64, _D, C.c_int64, _I32, _I32, _F]),
    "viso_ba_set_problem_batched": (C.c_int32, [_P, C.c_int64, _I64, _I64, _I64, _D, _D, _I32, _I32, _F]),
    "viso_ba_set_pose_image": (C.c_int32, [_P, C.c_int64, C.POINTER(C.c_uint8), C.c_int32, C.c_int32, C.c_int32]),
    "viso_ba_add_photo_edge": (C.c_int32, [_P, C.c_int64, C.c_int64, C.c_int64]),
    "viso_ba_set_photo_problem": (C.c_int32, [_P, C.c_int64, _D, C.c_int64, _D, C.c_int64, _I32, _I32, _I32]),
    "viso_ba_set_pose_fixed": (C.c_int32, [_P, C.c_int64, C.c_int32]),
    "viso_ba_set_intrinsics": (C.c_int32, [_P, C.c_double, C.c_double, C.c_double, C.c_double]),
    "viso_ba_set_huber_delta": (C.c_int32, [_P, C.c_double]),
    "viso_ba_nccl_unique_id": (C.c_int32, [_P]),
    "viso_ba_shard_init": (C.c_int32, [_P, C.c_int32, C.c_int32, _P]),
    "viso_ba_shard_point_ranges": (C.c_int32, [C.c_int64, C.c_int64, _I32, C.c_int32, _I64]),
    "viso_ba_set_stream": (C.c_int32, [_P, _P]),
    "viso_ba_upload": (C.c_int32, [_P]),
    "viso_ba_optimize": (C.c_int32, [_P, C.c_int32, C.POINTER(IterStats), C.POINTER(Summary)]),
    "viso_ba_reset_state": (C.c_int32, [_P]),
    "viso_ba_optimize_resident": (C.c_int32, [_P, C.c_int32, C.POINTER(Summary)]),
    "viso_ba_download": (C.c_int32, [_P]),
    "viso_ba_num_points": (C.c_int64, [_P]),
    "viso_ba_num_poses": (C.c_int64, [_P]),
    "viso_ba_num_edges": (C.c_int64, [_P]),
    "viso_ba_num_problems": (C.c_int64, [_P]),
    "viso_ba_get_poses": (C.c_int32, [_P, C.c_int64, C.c_int64, _D]),
    "viso_ba_get_points": (C.c_int32, [_P, C.c_int64, C.c_int64, _D]),
    "viso_ba_get_problem_summaries": (C.c_int32, [_P, C.POINTER(Summary)]),
    "viso_ba_eval_cost": (C.c_int32, [_P, _D]),
    # device pointers as plain addresses (the caller owns the device memory; nothing is allocated by the call)
    "viso_ba_eval_cost_static": (C.c_int32, [_P, C.c_int64, _P, C.c_int64, _P, C.c_int64, _P, _P, _P, C.c_double, C.c_double,
                                             C.c_double, C.c_double, C.c_double, _P, C.c_char_p, C.c_int32]),
    "viso_ba_debug_reduced_system": (C.c_int32, [_P, C.c_double, C.c_int32, _D, _D, _I32]),
    "viso_ba_debug_host_build": (C.c_int32, [_P, C.c_int32, _D, C.POINTER(C.c_uint64)]),
    "viso_ba_kernel_launches": (C.c_int64, [_P]),
    "viso_ba_set_kernel_timing": (C.c_int32, [_P, C.c_int32]),
    "viso_ba_get_kernel_times": (C.c_int32, [_P, C.c_int32, C.POINTER(KernelTime), _I32]),
}
# include/viso_ba_synth.h, exported by libviso_synth.so (no CUDA)
SYNTH_SYMBOLS = {
    "viso_ba_synth_preset": (C.c_int32, [C.c_int32, C.c_int32, C.POINTER(SynthSpec)]),
    "viso_ba_synth_create": (C.c_int32, [C.POINTER(SynthSpec), C.POINTER(_P)]),
    "viso_ba_synth_destroy": (None, [_P]),
    "viso_ba_synth_num_edges": (C.c_int64, [_P]),
    "viso_ba_synth_fetch": (C.c_int32, [_P, _D, _D, _D, _D, _I32, _I32, _F]),
}

_lib = None
_synth_lib = None


def load_synth_library():
    """dlopen libviso_synth.so (host-only: needs neither a GPU nor the CUDA backend)."""
    global _synth_lib
    if _synth_lib is None:
        if not os.path.exists(SYNTH_LIB_PATH):
            raise ImportError("%s is missing: build it with `python -m viso_b200.build`" % SYNTH_LIB_PATH)
        lib = C.CDLL(SYNTH_LIB_PATH)
        for name, (res, args) in SYNTH_SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _synth_lib = lib
    return _synth_lib


def load_library(path=None):
    """dlopen libviso_ba.so and declare every prototype.  Raises if the library is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise ImportError(
            "%s is missing: build it with `python -m viso_b200.build` (nvcc, sm_100a). "
            "This backend has no CPU fallback." % p)
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def default_config():
    cfg = Config()
    load_library().viso_ba_default_config(C.byref(cfg))
    return cfg


def _dptr(a):
    return a.ctypes.data_as(_D)


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


class BundleAdjuster:
    """Python mirror of the reference class (bundle_adjuster.h:108-187) over the C ABI.

    Reference-named methods: AddPoint, AddPose, AddEdge (bundle_adjuster.h:169-173).
    Completed interface: SetIntrinsics, SetPoseFixed, SetHuberDelta, Optimize, GetPose,
    GetPoint, Clear.  numpy bulk methods (lower-case) serve the large configurations.
    """

    def __init__(self, max_points=1000, max_poses=1000, max_edges=1000, **cfg_overrides):
        # defaults: the reference pools, bundle_adjuster.h:182-184
        self._lib = load_library()
        cfg = default_config()
        cfg.max_points, cfg.max_poses, cfg.max_edges = int(max_points), int(max_poses), int(max_edges)
        for k, v in cfg_overrides.items():
            if not hasattr(cfg, k):
                raise TypeError("unknown config field %r" % k)
            setattr(cfg, k, v)
        self._h = _P()
        st = self._lib.viso_ba_create(C.byref(cfg), C.byref(self._h))
        if st != OK:
            raise VisoBaError(st, "viso_ba_create failed")
        self.cfg = cfg

    # ---- lifetime -------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.viso_ba_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, st):
        if st != OK:
            raise VisoBaError(st, self._lib.viso_ba_last_error(self._h).decode())

    # ---- the reference's three methods ----------------------------------------------------
    def AddPoint(self, xyz):
        a = _c(xyz, np.float64).reshape(3)
        r = self._lib.viso_ba_add_point(self._h, _dptr(a))
        if r < 0:
            self._check(int(-r))
        return int(r)

    def AddPose(self, q_xyzw_t):
        a = _c(q_xyzw_t, np.float64).reshape(7)
        r = self._lib.viso_ba_add_pose(self._h, _dptr(a))
        if r < 0:
            self._check(int(-r))
        return int(r)

    def AddEdge(self, point, pose, u, v):
        self._check(self._lib.viso_ba_add_edge(self._h, int(point), int(pose), float(u), float(v)))

    # ---- the reference adjuster's own edge: EdgeDirectProjection (bundle_adjuster.h:58-106) ----
    def SetPoseImage(self, pose, image):
        """level-0 image (uint8, rows x cols) of the keyframe of pose vertex `pose`"""
        im = np.ascontiguousarray(image, dtype=np.uint8)
        self._check(self._lib.viso_ba_set_pose_image(self._h, int(pose), im.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                     im.shape[0], im.shape[1], im.strides[0]))

    def AddPhotoEdge(self, point, src_pose, dst_pose):
        self._check(self._lib.viso_ba_add_photo_edge(self._h, int(point), int(src_pose), int(dst_pose)))

    def set_photo_problem(self, poses7, points3, edge_point, edge_src, edge_dst, images=None):
        poses7, points3 = _c(poses7, np.float64), _c(points3, np.float64)
        el, es, ed = (_c(a, np.int32) for a in (edge_point, edge_src, edge_dst))
        self._check(self._lib.viso_ba_set_photo_problem(
            self._h, poses7.size // 7, _dptr(poses7), points3.size // 3, _dptr(points3), el.size,
            el.ctypes.data_as(_I32), es.ctypes.data_as(_I32), ed.ctypes.data_as(_I32)))
        for k, im in enumerate(images or []):
            self.SetPoseImage(k, im)

    # ---- completed interface --------------------------------------------------------------
    def SetIntrinsics(self, fx, fy, cx, cy):
        self._check(self._lib.viso_ba_set_intrinsics(self._h, fx, fy, cx, cy))

    def SetPoseFixed(self, pose, fixed=True):
        self._check(self._lib.viso_ba_set_pose_fixed(self._h, int(pose), 1 if fixed else 0))

    def SetHuberDelta(self, delta):
        self._check(self._lib.viso_ba_set_huber_delta(self._h, float(delta)))

    def Clear(self):
        self._check(self._lib.viso_ba_clear(self._h))

    def Optimize(self, iterations, with_stats=False):
        """LM for `iterations` outer iterations (H2D if needed, device loop, D2H).  Returns the
        Summary, or (Summary, [IterStats]) when with_stats."""
        s = Summary()
        stats = (IterStats * max(1, int(iterations)))() if with_stats else None
        self._check(self._lib.viso_ba_optimize(self._h, int(iterations), stats, C.byref(s)))
        if with_stats:
            return s, list(stats)[: s.iterations]
        return s

    def GetPose(self, i):
        out = np.empty(7)
        self._check(self._lib.viso_ba_get_poses(self._h, int(i), 1, _dptr(out)))
        return out

    def GetPoint(self, i):
        out = np.empty(3)
        self._check(self._lib.viso_ba_get_points(self._h, int(i), 1, _dptr(out)))
        return out

    # ---- numpy bulk interface -------------------------------------------------------------
    def set_problem(self, poses7, points3, edge_pose, edge_point, edge_uv):
        poses7, points3 = _c(poses7, np.float64), _c(points3, np.float64)
        edge_pose, edge_point = _c(edge_pose, np.int32), _c(edge_point, np.int32)
        edge_uv = _c(edge_uv, np.float32)
        self._check(self._lib.viso_ba_set_problem(
            self._h, poses7.size // 7, _dptr(poses7), points3.size // 3, _dptr(points3), edge_pose.size,
            edge_pose.ctypes.data_as(_I32), edge_point.ctypes.data_as(_I32), edge_uv.ctypes.data_as(_F)))

    def set_problem_batched(self, pose_off, point_off, edge_off, poses7, points3, edge_pose, edge_point, edge_uv):
        pose_off, point_off, edge_off = (_c(x, np.int64) for x in (pose_off, point_off, edge_off))
        poses7, points3 = _c(poses7, np.float64), _c(points3, np.float64)
        edge_pose, edge_point = _c(edge_pose, np.int32), _c(edge_point, np.int32)
        edge_uv = _c(edge_uv, np.float32)
        self._check(self._lib.viso_ba_set_problem_batched(
            self._h, pose_off.size - 1, pose_off.ctypes.data_as(_I64), point_off.ctypes.data_as(_I64),
            edge_off.ctypes.data_as(_I64), _dptr(poses7), _dptr(points3), edge_pose.ctypes.data_as(_I32),
            edge_point.ctypes.data_as(_I32), edge_uv.ctypes.data_as(_F)))

    def upload(self):
        self._check(self._lib.viso_ba_upload(self._h))

    def reset_state(self):
        self._check(self._lib.viso_ba_reset_state(self._h))

    def optimize_resident(self, iterations):
        s = Summary()
        self._check(self._lib.viso_ba_optimize_resident(self._h, int(iterations), C.byref(s)))
        return s

    def download(self):
        """device state -> host copy (what Optimize ends with); pairs with optimize_resident"""
        self._check(self._lib.viso_ba_download(self._h))

    @property
    def num_poses(self):
        return int(self._lib.viso_ba_num_poses(self._h))

    @property
    def num_points(self):
        return int(self._lib.viso_ba_num_points(self._h))

    @property
    def num_edges(self):
        return int(self._lib.viso_ba_num_edges(self._h))

    @property
    def num_problems(self):
        return int(self._lib.viso_ba_num_problems(self._h))

    def poses(self):
        out = np.empty((self.num_poses, 7))
        self._check(self._lib.viso_ba_get_poses(self._h, 0, out.shape[0], _dptr(out)))
        return out

    def points(self):
        out = np.empty((self.num_points, 3))
        self._check(self._lib.viso_ba_get_points(self._h, 0, out.shape[0], _dptr(out)))
        return out

    def problem_summaries(self):
        out = (Summary * self.num_problems)()
        self._check(self._lib.viso_ba_get_problem_summaries(self._h, out))
        return list(out)

    def eval_cost(self):
        c = C.c_double()
        self._check(self._lib.viso_ba_eval_cost(self._h, C.byref(c)))
        return c.value

    def reduced_system(self, lam):
        n = C.c_int32()
        self._check(self._lib.viso_ba_debug_reduced_system(self._h, float(lam), 0, None, None, C.byref(n)))
        S, bs = np.empty((n.value, n.value)), np.empty(n.value)
        self._check(self._lib.viso_ba_debug_reduced_system(self._h, float(lam), n.value, _dptr(S), _dptr(bs), C.byref(n)))
        return S, bs

    def host_build(self, threads=0):
        """host part of the general structure build alone (no GPU): (milliseconds, checksum of the tables)"""
        ms, ck = C.c_double(), C.c_uint64()
        self._check(self._lib.viso_ba_debug_host_build(self._h, int(threads), C.byref(ms), C.byref(ck)))
        return ms.value, ck.value

    def kernel_launches(self):
        return int(self._lib.viso_ba_kernel_launches(self._h))

    def set_kernel_timing(self, on):
        self._check(self._lib.viso_ba_set_kernel_timing(self._h, 1 if on else 0))

    def kernel_times(self):
        n = C.c_int32()
        self._check(self._lib.viso_ba_get_kernel_times(self._h, 0, None, C.byref(n)))
        arr = (KernelTime * max(1, n.value))()
        self._check(self._lib.viso_ba_get_kernel_times(self._h, n.value, arr, C.byref(n)))
        return {arr[i].name.decode(): (int(arr[i].launches), float(arr[i].device_ms)) for i in range(n.value)}

    def set_stream(self, cuda_stream_ptr):
        self._check(self._lib.viso_ba_set_stream(self._h, _P(cuda_stream_ptr)))

    def shard_init(self, rank, n_ranks, nccl_id):
        buf = C.create_string_buffer(bytes(nccl_id), 128)
        self._check(self._lib.viso_ba_shard_init(self._h, int(rank), int(n_ranks), C.cast(buf, _P)))


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    st = load_library().viso_ba_nccl_unique_id(C.cast(buf, _P))
    if st != OK:
        raise VisoBaError(st, "ncclGetUniqueId failed (libnccl.so.2 not loadable?)")
    return buf.raw


def shard_point_ranges(n_points, edge_point, n_ranks):
    """the contiguous point ranges a sharded handle assigns to the ranks (host-only)"""
    edge_point = _c(edge_point, np.int32)
    out = np.zeros(n_ranks + 1, np.int64)
    st = load_library().viso_ba_shard_point_ranges(int(n_points), edge_point.size, edge_point.ctypes.data_as(_I32),
                                                   int(n_ranks), out.ctypes.data_as(_I64))
    if st != OK:
        raise VisoBaError(st, "viso_ba_shard_point_ranges failed")
    return out


# ---- synthetic problems; host-only, needs no GPU ------------------------------------
def synth_preset(config_id, window=0):
    sp = SynthSpec()
    st = load_synth_library().viso_ba_synth_preset(int(config_id), int(window), C.byref(sp))
    if st != OK:
        raise VisoBaError(st, "unknown synthetic preset %r" % (config_id,))
    return sp


def synth_problem(spec):
    """Generate one problem; returns a dict of numpy arrays."""
    lib = load_synth_library()
    s = _P()
    st = lib.viso_ba_synth_create(C.byref(spec), C.byref(s))
    if st != OK:
        raise VisoBaError(st, "viso_ba_synth_create failed")
    try:
        ne = int(lib.viso_ba_synth_num_edges(s))
        out = {
            "poses_init": np.empty((spec.n_poses, 7)), "poses_true": np.empty((spec.n_poses, 7)),
            "points_init": np.empty((spec.n_points, 3)), "points_true": np.empty((spec.n_points, 3)),
            "edge_pose": np.empty(ne, np.int32), "edge_point": np.empty(ne, np.int32),
            "edge_uv": np.empty((ne, 2), np.float32),
        }
        st = lib.viso_ba_synth_fetch(
            s, _dptr(out["poses_init"]), _dptr(out["poses_true"]), _dptr(out["points_init"]),
            _dptr(out["points_true"]), out["edge_pose"].ctypes.data_as(_I32),
            out["edge_point"].ctypes.data_as(_I32), out["edge_uv"].ctypes.data_as(_F))
        if st != OK:
            raise VisoBaError(st, "viso_ba_synth_fetch failed")
    finally:
        lib.viso_ba_synth_destroy(s)
    out["K"] = (spec.fx, spec.fy, spec.cx, spec.cy)
    out["huber_delta"] = 2.0 if spec.outlier_frac > 0 else 0.0
    return out


def synth_config(config_id, window=0, **overrides):
    sp = synth_preset(config_id, window)
    for k, v in overrides.items():
        setattr(sp, k, v)
    return synth_problem(sp)


def batch_problems(problems):
    """Concatenate independent problems into the arrays viso_ba_set_problem_batched takes."""
    pose_off = np.zeros(len(problems) + 1, np.int64)
    point_off = np.zeros(len(problems) + 1, np.int64)
    edge_off = np.zeros(len(problems) + 1, np.int64)
    for i, p in enumerate(problems):
        pose_off[i + 1] = pose_off[i] + len(p["poses_init"])
        point_off[i + 1] = point_off[i] + len(p["points_init"])
        edge_off[i + 1] = edge_off[i] + len(p["edge_pose"])
    return {
        "pose_off": pose_off, "point_off": point_off, "edge_off": edge_off,
        "poses_init": np.concatenate([p["poses_init"] for p in problems]),
        "points_init": np.concatenate([p["points_init"] for p in problems]),
        "edge_pose": np.concatenate([p["edge_pose"] for p in problems]),
        "edge_point": np.concatenate([p["edge_point"] for p in problems]),
        "edge_uv": np.concatenate([p["edge_uv"] for p in problems]),
    }
