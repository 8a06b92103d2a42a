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


def run_batched(nwin, iters=10, check=4):
    t = time.time()
    probs = [vb.synth_config(5, w) for w in range(nwin)]
    b = vb.batch_problems(probs)
    print("batched", nwin, "windows; gen", time.time() - t, "s; edges", len(b["edge_pose"]))
    ba = vb.BundleAdjuster(len(b["points_init"]), len(b["poses_init"]), len(b["edge_pose"]))
    t = time.time()
    ba.set_problem_batched(b["pose_off"], b["point_off"], b["edge_off"], b["poses_init"], b["points_init"], b["edge_pose"], b["edge_point"], b["edge_uv"])
    t1 = time.time(); ba.upload(); t2 = time.time()
    print("  set_problem", t1 - t, "upload", t2 - t1)
    s = ba.Optimize(iters)
    print("  optimize", s.as_dict(), "wall", time.time() - t2)
    sums = ba.problem_summaries()
    poses, pts = ba.poses(), ba.points()
    c = orc.cfg(threads=8)
    for w in range(min(check, nwin)):
        p = probs[w]
        po, pt, fc, it, so = orc.optimize(c, p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"], iters)
        gp = poses[b["pose_off"][w]:b["pose_off"][w + 1]]
        gq = pts[b["point_off"][w]:b["point_off"][w + 1]]
        print("   win", w, "cost rel", abs(sums[w].final_cost - fc) / fc, "pose", np.abs(gp - po).max(), "pt", np.abs(gq - pt).max(), "trials", sums[w].trials, sum(x.trials for x in so))
    ba.reset_state(); ba.set_kernel_timing(True)
    s2 = ba.optimize_resident(iters)
    kt = ba.kernel_times()
    print("  resident timed", s2.device_ms, "ms;", {k: (v[0], round(v[1], 3)) for k, v in kt.items()})
    ba.set_kernel_timing(False)
    for _ in range(3):
        ba.reset_state(); s3 = ba.optimize_resident(iters)
    ne = len(b["edge_pose"])
    print("  resident graph", s3.device_ms, "ms ->", ne * s3.trials / nwin / (s3.device_ms * 1e-3) / 1e9, "G obs/s; trials", s3.trials)


if __name__ == "__main__" and os.environ.get("BATCH"):
    run_batched(int(os.environ["BATCH"]))
