"""GPU box: kernel time of the direct pose estimator against the number of map points (fixed cost vs per-point cost)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
from viso_b200 import direct
from viso_b200.direct_synth import synth_scene
sc = synth_scene(seed=11, n_points=8000)
for n in (16, 128, 500, 2000, 8000):
    est = direct.DirectPoseEstimator(max_points=8000, reference_termination=0)
    est.set_pyramid(0, sc["ref_pyr"]); est.set_pyramid(1, sc["cur_pyr"]); est.set_points(sc["points"][:n])
    ms = []
    for i in range(12):
        pose, st, m = est.DirectPoseEstimationMultiLayer(sc["identity"])
        ms.append(m)
    its = sum(s.iterations for s in st[0])
    print("points %5d  iterations %3d  kernel %.3f ms  %.1f us/iteration" % (n, its, np.median(ms), 1e3 * np.median(ms) / its))
