"""Synthetic input for the direct pose estimator (numpy; input generation only, not a compute path).

Scene: a textured plane at depth 1 in the last frame's camera (viso normalises the mean scene depth
to 1, reference src/viso.cpp:622-637), texture = an analytic sum of sinusoids, so the current frame
is rendered exactly (no resampling): every current pixel's ray is intersected with the plane and the
texture evaluated at its projection into the last frame.  Intrinsics and image size are the
reference's (src/main.cpp:14-17, :73).  Pyramids follow Keyframe's constructor (include/keyframe.h:
29-45): cv::pyrDown to half size, 4 levels; if cv2 is missing a 2x2 box filter stands in.
"""
import numpy as np

FX, FY, CX, CY = 517.3, 516.5, 325.1, 249.7
W, H = 640, 480


def quat_to_R(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def pose_from_rotvec(rv, t):
    rv = np.asarray(rv, dtype=np.float64)
    th = np.linalg.norm(rv)
    if th < 1e-12:
        q = np.array([0.5 * rv[0], 0.5 * rv[1], 0.5 * rv[2], 1.0])
    else:
        q = np.concatenate([np.sin(0.5 * th) * rv / th, [np.cos(0.5 * th)]])
    q /= np.linalg.norm(q)
    return np.concatenate([q, np.asarray(t, dtype=np.float64)])


def _texture(rng, n_waves=24):
    freq = rng.uniform(0.02, 0.35, size=(n_waves, 2)) * rng.choice([-1.0, 1.0], size=(n_waves, 2))
    phase = rng.uniform(0, 2 * np.pi, size=n_waves)
    amp = rng.uniform(6.0, 22.0, size=n_waves)

    def T(u, v):
        s = np.full(np.shape(u), 128.0)
        for (a, b), p, m in zip(freq, phase, amp):
            s = s + m * np.sin(a * u + b * v + p)
        return s
    return T


def pyramid(img, levels=4):
    pyr = [np.ascontiguousarray(img)]
    try:
        import cv2
        for _ in range(1, levels):
            prev = pyr[-1]
            pyr.append(np.ascontiguousarray(cv2.pyrDown(prev, dstsize=(prev.shape[1] // 2, prev.shape[0] // 2))))
    except ImportError:
        for _ in range(1, levels):
            p = pyr[-1].astype(np.float64)
            h, w = (p.shape[0] // 2) * 2, (p.shape[1] // 2) * 2
            q = 0.25 * (p[0:h:2, 0:w:2] + p[1:h:2, 0:w:2] + p[0:h:2, 1:w:2] + p[1:h:2, 1:w:2])
            pyr.append(np.ascontiguousarray(np.clip(np.rint(q), 0, 255).astype(np.uint8)))
    return pyr


def synth_scene(seed=1, n_points=2000, rotvec=(0.004, -0.006, 0.003), trans=(0.012, -0.008, 0.01),
                init_noise=0.0, width=W, height=H):
    """Returns dict: ref_pyr, cur_pyr (lists of 4 uint8 images), points [n][3] (world = last frame's
    camera frame), ref_pose [7] (identity), true_pose [7] (current frame), init_pose [7]."""
    rng = np.random.default_rng(seed)
    T = _texture(rng)
    uu, vv = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    ref = np.clip(np.rint(T(uu, vv)), 0, 255).astype(np.uint8)
    true_pose = pose_from_rotvec(rotvec, trans)
    R, t = quat_to_R(true_pose[:4]), true_pose[4:]
    # current pixel -> ray -> plane z_ref = 1 -> last frame's pixel
    d = np.stack([(uu - CX) / FX, (vv - CY) / FY, np.ones_like(uu)], axis=-1)
    Rt_d = d @ R  # rows: R^T d
    Rt_t = R.T @ t
    s = (1.0 + Rt_t[2]) / Rt_d[..., 2]
    Xr = Rt_d * s[..., None] - Rt_t
    ur, vr = FX * Xr[..., 0] / Xr[..., 2] + CX, FY * Xr[..., 1] / Xr[..., 2] + CY
    cur = np.clip(np.rint(T(ur, vr)), 0, 255).astype(np.uint8)
    # map points on the plane, spread over the image with a margin
    pu = rng.uniform(30, width - 30, size=n_points)
    pv = rng.uniform(30, height - 30, size=n_points)
    pts = np.stack([(pu - CX) / FX, (pv - CY) / FY, np.ones(n_points)], axis=-1)
    init = true_pose.copy() if init_noise == 0.0 else None
    if init is None:
        init = pose_from_rotvec(np.asarray(rotvec) + rng.normal(0, init_noise, 3), np.asarray(trans) + rng.normal(0, init_noise, 3))
    return {"ref_pyr": pyramid(ref), "cur_pyr": pyramid(cur), "points": pts,
            "ref_pose": np.array([0, 0, 0, 1.0, 0, 0, 0]), "true_pose": true_pose, "init_pose": init,
            "identity": np.array([0, 0, 0, 1.0, 0, 0, 0])}


def render_plane(T, pose, width=W, height=H):
    """The textured plane z = 1 of the world frame seen from `pose` (world -> camera), rendered exactly."""
    uu, vv = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    R, t = quat_to_R(pose[:4]), pose[4:]
    d = np.stack([(uu - CX) / FX, (vv - CY) / FY, np.ones_like(uu)], axis=-1)
    Rt_d = d @ R
    Rt_t = R.T @ t
    s = (1.0 + Rt_t[2]) / Rt_d[..., 2]
    Xw = Rt_d * s[..., None] - Rt_t
    return np.clip(np.rint(T(FX * Xw[..., 0] / Xw[..., 2] + CX, FY * Xw[..., 1] / Xw[..., 2] + CY)), 0, 255).astype(np.uint8)


def project(pose, P):
    """Keyframe::Project at level 0 for an array of points"""
    Pc = P @ quat_to_R(pose[:4]).T + pose[4:]
    return np.stack([FX * Pc[:, 0] / Pc[:, 2] + CX, FY * Pc[:, 1] / Pc[:, 2] + CY], axis=1)


def synth_sequence(seed=1, n_frames=5, n_points=1500, step_rot=(0.003, -0.004, 0.002), step_trans=(0.010, -0.006, 0.008)):
    """A short camera trajectory over the textured plane: frame 0 at the identity (viso keeps frame 0 there,
    include/keyframe.h:34-35), frame k at k steps of a constant twist.  Returns images, true poses, map points."""
    rng = np.random.default_rng(seed)
    T = _texture(rng)
    poses = [pose_from_rotvec(np.asarray(step_rot) * k, np.asarray(step_trans) * k) for k in range(n_frames)]
    images = [render_plane(T, p) for p in poses]
    pu = rng.uniform(60, W - 60, size=n_points)
    pv = rng.uniform(60, H - 60, size=n_points)
    pts = np.stack([(pu - CX) / FX, (pv - CY) / FY, np.ones(n_points)], axis=-1)
    return {"images": images, "poses": np.stack(poses), "points": pts}


def synth_photo_window(seed=1, n_frames=4, n_points=300, width=320, height=240, pose_noise=1e-3, depth_noise=0.005,
                       step_rot=(0.004, -0.005, 0.003), step_trans=(0.03, -0.02, 0.015)):
    """A small window for bundle adjustment over the reference adjuster's photometric edge (reference
    include/bundle_adjuster.h:58-106): n_frames keyframes over the textured plane, map points on it, and one edge
    (point, source keyframe, pose vertex) from each point's host keyframe (0 or 1) to every other keyframe.  The
    initial poses of keyframes >= 2 and the point depths are perturbed; keyframe 0 stays at the identity and keyframe 1
    at its true pose (with both fixed the scale of the window is pinned; photometric edges alone leave it free).
    Intrinsics are the reference's scaled to the image size."""
    rng = np.random.default_rng(seed)
    T = _texture(rng)
    sx, sy = width / W, height / H
    K = (FX * sx, FY * sy, CX * sx, CY * sy)
    true = [pose_from_rotvec(np.asarray(step_rot) * k, np.asarray(step_trans) * k) for k in range(n_frames)]

    def render(pose):
        uu, vv = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
        R, t = quat_to_R(pose[:4]), pose[4:]
        d = np.stack([(uu - K[2]) / K[0], (vv - K[3]) / K[1], np.ones_like(uu)], axis=-1)
        Rt_d, Rt_t = d @ R, R.T @ t
        s = (1.0 + Rt_t[2]) / Rt_d[..., 2]
        Xw = Rt_d * s[..., None] - Rt_t
        # the texture lives on the plane z = 1 of the world frame, parametrised by frame 0's full-size pixel grid
        return np.clip(np.rint(T(FX * Xw[..., 0] / Xw[..., 2] + CX, FY * Xw[..., 1] / Xw[..., 2] + CY)), 0, 255).astype(np.uint8)
    images = [render(p) for p in true]
    pu = rng.uniform(0.2 * width, 0.8 * width, size=n_points)
    pv = rng.uniform(0.2 * height, 0.8 * height, size=n_points)
    pts_true = np.stack([(pu - K[2]) / K[0], (pv - K[3]) / K[1], np.ones(n_points)], axis=-1)
    poses_init = np.stack(true).copy()
    for k in range(2, n_frames):
        poses_init[k] = pose_from_rotvec(np.asarray(step_rot) * k + rng.normal(0, pose_noise, 3),
                                         np.asarray(step_trans) * k + rng.normal(0, pose_noise, 3))
    pts_init = pts_true * (1.0 + rng.normal(0, depth_noise, size=(n_points, 1)))
    e_point, e_src, e_dst = [], [], []
    for p in range(n_points):
        host = p % 2
        for k in range(n_frames):
            if k != host:
                e_point.append(p), e_src.append(host), e_dst.append(k)
    return {"K": K, "images": images, "poses_true": np.stack(true), "points_true": pts_true, "poses_init": poses_init,
            "points_init": pts_init, "edge_point": np.array(e_point, np.int32), "edge_src": np.array(e_src, np.int32),
            "edge_dst": np.array(e_dst, np.int32)}
