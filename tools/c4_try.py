#!/usr/bin/env python
"""BASELINE configs[3] on one GPU: 2k cameras, 1M points, 10M observations (dense reduced system, n = 11994)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import viso_b200 as vb

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
t0 = time.time()
p = vb.synth_config(4)
t1 = time.time()
ba = vb.BundleAdjuster(len(p["points_init"]), len(p["poses_init"]), len(p["edge_pose"]), device=0)
ba.set_problem(p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
ba.upload()
t2 = time.time()
ba.set_kernel_timing(True)
s = ba.optimize_resident(iters)
kt = ba.kernel_times()
t3 = time.time()
out = {"gen_s": t1 - t0, "upload_s": t2 - t1, "opt_wall_s": t3 - t2, "device_ms": s.device_ms, "iterations": s.iterations, "trials": s.trials,
       "initial_cost": s.initial_cost, "final_cost": s.final_cost, "obs": len(p["edge_pose"]),
       "kernels": {k: (n, round(ms, 3)) for k, (n, ms) in kt.items()}}
print(json.dumps(out))
