"""Generate tests/golden/*.npz with the independent numpy restatement (tests/numpy_ba.py).

The reference adjuster cannot be compiled or imported, so these vectors do
NOT pin parity with the reference; they pin the C oracle and the CUDA path against an
independently written fp64 implementation of the same published algorithm.  Inputs come
from the repo's deterministic synthetic generator.  Run:  python tools/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy_ba  # noqa: E402
import viso_b200 as vb  # noqa: E402

CASES = {
    # name: (synth overrides, huber delta, LM iterations, lambda for the reduced system)
    "tiny_l2": (dict(n_poses=4, n_points=40, target_edges=130, seed=0x601D0001), 0.0, 6, 3.5),
    "tiny_huber": (dict(n_poses=5, n_points=60, target_edges=240, seed=0x601D0002, outlier_frac=0.08), 2.0, 6, 0.25),
    "tiny_dup_fixed2": (dict(n_poses=6, n_points=50, target_edges=220, seed=0x601D0003), 0.0, 5, 10.0),
    # far start + almost no damping: the first steps overshoot and LM has to reject trials
    "tiny_reject": (dict(n_poses=5, n_points=50, target_edges=200, seed=0x601D0006, point_sigma=0.25,
                         pose_rot_sigma=0.08, pose_trans_sigma=0.15), 0.0, 8, 1e-3),
}
TAU = {"tiny_reject": 1e-8}


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, (ov, delta, iters, lam) in CASES.items():
        p = vb.synth_config(1, **ov)
        ep, el, uv = p["edge_pose"].copy(), p["edge_point"].copy(), p["edge_uv"].copy()
        fixed = [0]
        if name == "tiny_dup_fixed2":  # duplicate (point, pose) edges and a second fixed pose
            ep = np.concatenate([ep, ep[:7]])
            el = np.concatenate([el, el[:7]])
            uv = np.concatenate([uv, uv[:7] + np.float32(0.5)])
            fixed = [0, 3]
        prob = numpy_ba.Problem(p["K"], p["poses_init"], p["points_init"], ep, el, uv, delta, fixed)
        cost0 = prob.cost()
        S, bs = prob.reduced_system(lam)
        hist = prob.lm(iters, tau=TAU.get(name, 1e-5))
        np.savez(
            os.path.join(out_dir, name + ".npz"),
            K=np.array(p["K"]), huber_delta=delta, tau=TAU.get(name, 1e-5), fixed=np.array(fixed), iterations=iters, lam=lam,
            poses_init=p["poses_init"], points_init=p["points_init"], edge_pose=ep, edge_point=el, edge_uv=uv,
            cost0=cost0, S=S, bs=bs, hist=np.array(hist), poses_final=prob.poses7(), points_final=prob.points3())
        print(name, "edges", len(ep), "cost0 %.6f" % cost0, "final %.9f" % hist[-1][0], "trials", [h[2] for h in hist])


if __name__ == "__main__":
    main()
