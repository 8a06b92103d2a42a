"""Where the end-to-end time of a LARGE problem goes (general path, BASELINE configs 3 and 4) on one GPU: set_problem (host
staging), upload (host structure build + H2D + device tables; verbose prints the phases of the host build), LM, read-back."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import viso_b200 as vb

for cid in (3, 4):
    p = vb.synth_config(cid)
    ba = vb.BundleAdjuster(len(p["points_init"]), len(p["poses_init"]), len(p["edge_pose"]), device=0, verbose=1)
    for rep in range(2):
        t0 = time.perf_counter()
        ba.set_problem(p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
        t1 = time.perf_counter()
        ba.upload()
        t2 = time.perf_counter()
        s = ba.optimize_resident(10)
        t3 = time.perf_counter()
        ba.download()
        t4 = time.perf_counter()
        print("config %d rep %d: set_problem %.1f ms, upload %.1f ms, LM %.1f ms (device %.1f), download %.1f ms" % (
            cid, rep, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), s.device_ms, 1e3 * (t4 - t3)), flush=True)
    ba.close()
