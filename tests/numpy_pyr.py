"""numpy statement of cv::pyrDown for CV_8UC1 as Keyframe's constructor calls it
(reference include/keyframe.h:40-45: cv::pyrDown(src, dst, Size(cols * 0.5, rows * 0.5))).
OpenCV (the reference's dependency, not under /root/reference; imgproc/pyramids.cpp): separable
5-tap kernel [1 4 6 4 1] / 16 in each direction, centred on (2x, 2y), BORDER_REFLECT_101, integer
arithmetic, result (sum + 128) >> 8.  Pinned by tests/golden/pyrdown_cv2.npz, which holds outputs
of cv2.pyrDown itself (tools/make_golden_pyr.py)."""
import numpy as np


def reflect101(i, n):
    i = np.abs(i)
    return np.where(i >= n, 2 * (n - 1) - i, i)


def pyr_down(src):
    src = np.asarray(src, dtype=np.uint8)
    rows, cols = src.shape
    drows, dcols = int(rows * 0.5), int(cols * 0.5)
    w = np.array([1, 4, 6, 4, 1], dtype=np.int64)
    s = src.astype(np.int64)
    xs = reflect101(2 * np.arange(dcols)[:, None] + np.arange(-2, 3)[None, :], cols)  # [dcols][5]
    hor = (s[:, xs] * w).sum(axis=2)                                                   # [rows][dcols]
    ys = reflect101(2 * np.arange(drows)[:, None] + np.arange(-2, 3)[None, :], rows)   # [drows][5]
    ver = (hor[ys, :] * w[None, :, None]).sum(axis=1)                                   # [drows][dcols]
    return ((ver + 128) >> 8).astype(np.uint8)


def pyramid(img, levels=4):
    pyr = [np.ascontiguousarray(img, dtype=np.uint8)]
    for _ in range(1, levels):
        pyr.append(pyr_down(pyr[-1]))
    return pyr
