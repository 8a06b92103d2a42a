"""Where the time of ONE window goes (BASELINE configs 1 and 2): per-kernel device times of the LM loop on both paths."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import viso_b200 as vb

for cid in (1, 2):
    p = vb.synth_config(cid)
    for path in (1, 2):
        ba = vb.BundleAdjuster(len(p["points_init"]), len(p["poses_init"]), len(p["edge_pose"]), device=0, path=path,
                               huber_delta=(1.0 if cid == 2 else 0.0))
        ba.set_problem(p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
        ba.upload()
        best = 1e9
        for _ in range(5):
            ba.reset_state()
            s = ba.optimize_resident(10)
            best = min(best, s.device_ms)
        ba.set_kernel_timing(True)
        ba.reset_state()
        s2 = ba.optimize_resident(10)
        kt = ba.kernel_times()
        print("config", cid, "path", path, "poses", len(p["poses_init"]), "edges", len(p["edge_pose"]), "LM loop %.3f ms (%d trials), final cost %.6f" % (best, s.trials, s.final_cost))
        print("   ", {k: (v[0], round(1e3 * v[1] / max(1, v[0]), 1)) for k, v in kt.items()})
