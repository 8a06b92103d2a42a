"""First GPU bring-up: C1 through the C ABI against the oracle, step by step."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import viso_b200 as vb
import oracle_lib as orc

def run(cfg_id, iters=10, **kw):
    p = vb.synth_config(cfg_id)
    ne = len(p["edge_pose"])
    hub = p["huber_delta"]
    ba = vb.BundleAdjuster(len(p["points_init"]), len(p["poses_init"]), ne, huber_delta=hub, **kw)
    ba.set_problem(p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
    c = orc.cfg(huber_delta=hub, threads=8)
    c0 = orc.cost(c, p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
    g0 = ba.eval_cost()
    print("cfg", cfg_id, "edges", ne, "cost oracle", c0, "gpu", g0, "rel", abs(c0 - g0) / c0)
    if len(p["poses_init"]) <= 30:
        S, bs = ba.reduced_system(1.0)
        So, bso, md = orc.reduced_system(c, p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"], 1.0)
        print("  S rel err", np.abs(S - So).max() / np.abs(So).max(), "bs rel err", np.abs(bs - bso).max() / np.abs(bso).max())
    t = time.time()
    s, st = ba.Optimize(iters, with_stats=True)
    tg = time.time() - t
    t = time.time()
    po, pt, fc, it, so = orc.optimize(c, p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"], iters)
    to = time.time() - t
    print("  gpu", s.as_dict(), "wall", tg)
    print("  oracle final", fc, "iters", it, "wall", to)
    for a, b in zip(st, so):
        print("   it gpu %.9e -> %.9e lam %.3e tr %d acc %d | orc %.9e -> %.9e lam %.3e tr %d acc %d" % (
            a.cost_before, a.cost_after, a.lambda_, a.trials, a.accepted, b.cost_before, b.cost_after, b.lambda_, b.trials, b.accepted))
    print("  cost rel", abs(s.final_cost - fc) / fc, "pose max", np.abs(ba.poses() - po).max(), "point max", np.abs(ba.points() - pt).max())
    ba.reset_state()
    ba.set_kernel_timing(True)
    s2 = ba.optimize_resident(iters)
    print("  resident", s2.device_ms, "ms; kernels", ba.kernel_times())
    ba.set_kernel_timing(False)
    for _ in range(3):
        ba.reset_state(); s3 = ba.optimize_resident(iters)
    print("  resident graph", s3.device_ms, "ms", "launches", ba.kernel_launches())

if __name__ == "__main__":
    for c in [int(x) for x in sys.argv[1:]] or [1, 2]:
        run(c)
