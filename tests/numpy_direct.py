"""Independent numpy statement of one linearisation of viso's direct pose estimation
(reference src/viso.cpp:688-733), vectorised over points and patch pixels.  Written from the
reference text, not from oracle/direct_oracle.c, so that a shared misreading shows up as a
disagreement."""
import numpy as np

from viso_b200.direct_synth import quat_to_R

SCALES = (1.0, 0.5, 0.25, 0.125)


def bilinear(img, x, y):
    """Keyframe::GetPixelValue, include/keyframe.h:50-57"""
    xi, yi = np.trunc(x).astype(np.int64), np.trunc(y).astype(np.int64)
    xx, yy = x - np.floor(x), y - np.floor(y)
    f = img.astype(np.float64).ravel()  # the reference indexes the flat cv::Mat buffer: data[int(y) * step + int(x)]
    step = img.shape[1]

    def tap(i):
        return f[np.clip(i, 0, f.size - 1)]  # past the buffer (undefined in the reference): clamped
    base = yi * step + xi
    return ((1 - xx) * (1 - yy) * tap(base) + xx * (1 - yy) * tap(base + 1)
            + (1 - xx) * yy * tap(base + step) + xx * yy * tap(base + step + 1))


def project(K, pose, P, scale):
    """Keyframe::Project, include/keyframe.h:82-89"""
    fx, fy, cx, cy = K
    Pc = P @ quat_to_R(pose[:4]).T + pose[4:]
    return scale * (Pc[:, 0] / Pc[:, 2] * fx + cx), scale * (Pc[:, 1] / Pc[:, 2] * fy + cy), Pc


def linearize(K, level, ref, cur, ref_pose, T21, points, hp=4):
    fx, fy, cx, cy = K
    scale = SCALES[level]
    ur, vr, _ = project(K, ref_pose, points, scale)
    uc, vc, Pc = project(K, T21, points, scale)

    def inside(img, u, v):
        return (u >= 0) & (u < img.shape[1]) & (v >= 0) & (v < img.shape[0])
    good = (inside(ref, ur - hp, vr - hp) & inside(ref, ur + hp, vr + hp)
            & inside(cur, uc - hp, vc - hp) & inside(cur, uc + hp, vc + hp))
    ur, vr, uc, vc, Pc = ur[good], vr[good], uc[good], vc[good], Pc[good]
    x, y, z = Pc[:, 0], Pc[:, 1], Pc[:, 2]
    sfx, sfy = fx * scale, fy * scale
    zero = np.zeros_like(x)
    # dPixeldXi, src/viso.cpp:640-658
    J0 = np.stack([sfx / z, zero, -sfx * x / z**2, -sfx * x * y / z**2, sfx + sfx * x * x / z**2, -sfx * y / z], axis=1)
    J1 = np.stack([zero, sfy / z, -sfy * y / z**2, -sfy - sfy * y * y / z**2, sfy * x * y / z**2, sfy * x / z], axis=1)
    d = np.arange(-hp, hp, dtype=np.float64)
    dx, dy = np.meshgrid(d, d, indexing="ij")
    dx, dy = dx.ravel()[None, :], dy.ravel()[None, :]
    U, V = uc[:, None] + dx, vc[:, None] + dy
    err = bilinear(ref, ur[:, None] + dx, vr[:, None] + dy) - bilinear(cur, U, V)
    gx = 0.5 * (bilinear(cur, U + 1, V) - bilinear(cur, U - 1, V))
    gy = 0.5 * (bilinear(cur, U, V + 1) - bilinear(cur, U, V - 1))
    J = -(gx[:, :, None] * J0[:, None, :] + gy[:, :, None] * J1[:, None, :])  # [n][64][6]
    J = J.reshape(-1, 6)
    e = err.ravel()
    return J.T @ J, -(J.T @ e), float(e @ e), int(good.sum())
