"""Independent numpy statement of the photometric edge of the reference adjuster as a residual WITH Jacobians
(reference include/bundle_adjuster.h:70-101 for the error; the Jacobian is image gradient, include/keyframe.h:59-64,
times dPixeldXi, src/viso.cpp:640-658, and times the projection Jacobians of both frames for the point).  Written
from the reference text, not from oracle/photo_ba_oracle.c; vectorised over the 16 samples of one edge."""
import numpy as np

from numpy_direct import bilinear
from viso_b200.direct_synth import quat_to_R


def _geom(K, pose, P):
    fx, fy, cx, cy = K
    R = quat_to_R(pose[:4])
    x, y, z = R @ P + pose[4:]
    # dPixeldXi, src/viso.cpp:653-655 (scale 1)
    B = np.array([[fx / z, 0, -fx * x / z**2, -fx * x * y / z**2, fx + fx * x * x / z**2, -fx * y / z],
                  [0, fy / z, -fy * y / z**2, -fy - fy * y * y / z**2, fy * x * y / z**2, fy * x / z]])
    A = B[:, :3] @ R  # d pixel / d world point
    return np.array([fx * x / z + cx, fy * y / z + cy]), B, A


def edge(K, src_img, tgt_img, src_pose, dst_pose, P, smooth=False):
    """-> (valid, r[16], Jp[16][6], Jl[16][3]).  smooth=True drops the float casts of the sample coordinates (so that the
    residual can be differenced numerically)."""
    uv_t, B, At = _geom(K, dst_pose, P)
    uv_s, _, As = _geom(K, src_pose, P)
    i, j = np.meshgrid(np.arange(-2, 2), np.arange(-2, 2))  # idx = 4 (j + 2) + (i + 2): j outer, i inner
    i, j = i.ravel().astype(np.float64), j.ravel().astype(np.float64)
    cast = (lambda a: a) if smooth else (lambda a: a.astype(np.float32).astype(np.float64))
    u1, v1, u2, v2 = cast(uv_t[0] + i), cast(uv_t[1] + j), cast(uv_s[0] + i), cast(uv_s[1] + j)

    def inside(img, u, v):
        return (u >= 0) & (u < img.shape[1]) & (v >= 0) & (v < img.shape[0])
    if not (inside(tgt_img, u1, v1) & inside(src_img, u2, v2)).all():
        return False, np.zeros(16), np.zeros((16, 6)), np.zeros((16, 3))
    r = bilinear(src_img, u2, v2) - bilinear(tgt_img, u1, v1)

    def grad(img, u, v):
        return np.stack([0.5 * (bilinear(img, u + 1, v) - bilinear(img, u - 1, v)),
                         0.5 * (bilinear(img, u, v + 1) - bilinear(img, u, v - 1))], axis=1)
    Gt, Gs = grad(tgt_img, u1, v1), grad(src_img, u2, v2)
    return True, r, -Gt @ B, Gs @ As - Gt @ At
