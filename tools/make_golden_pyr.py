"""Golden vectors for the device pyrDown: outputs of cv2.pyrDown itself — OpenCV is the library the
reference calls at include/keyframe.h:40-45 — on seeded inputs.  Run in the build container (cv2
4.13 is present there); writes tests/golden/pyrdown_cv2.npz."""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from viso_b200.direct_synth import synth_scene  # noqa: E402

rng = np.random.default_rng(0x5EED)
out = {"cv2_version": np.array(cv2.__version__)}
cases = {"rand_120x160": rng.integers(0, 256, size=(120, 160), dtype=np.uint8),
         "rand_61x81": rng.integers(0, 256, size=(61, 81), dtype=np.uint8),       # odd sizes: Size(cols * 0.5, rows * 0.5) truncates
         "rand_37x100": rng.integers(0, 256, size=(37, 100), dtype=np.uint8),
         "extremes_48x64": rng.choice(np.array([0, 255], dtype=np.uint8), size=(48, 64)),
         "texture_480x640": synth_scene(seed=4, n_points=4)["ref_pyr"][0]}
for name, img in cases.items():
    out[name + "_l0"] = img
    cur = img
    for level in range(1, 4):  # Keyframe: 4 levels, each cv::pyrDown of the previous to Size(cols * 0.5, rows * 0.5)
        cur = cv2.pyrDown(cur, dstsize=(int(cur.shape[1] * 0.5), int(cur.shape[0] * 0.5)))
        out["%s_l%d" % (name, level)] = cur
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "pyrdown_cv2.npz"), **out)
print("wrote", len(out) - 1, "arrays")
