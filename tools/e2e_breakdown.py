#!/usr/bin/env python
"""Where the end-to-end time of workload c5 goes: set_problem (host checks, staging, H2D of the edges),
upload (tables, structure build on the device), LM loop, read-back."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import viso_b200 as vb
W = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
probs = [vb.synth_config(5, w) for w in range(W)]
b = vb.batch_problems(probs)
host = {}
keep = []
for k, v in b.items():
    t = torch.from_numpy(np.ascontiguousarray(v)).pin_memory(); keep.append(t); host[k] = t.numpy()
ba = vb.BundleAdjuster(len(host["points_init"]), len(host["poses_init"]), len(host["edge_pose"]), device=0)
def setp():
    ba.set_problem_batched(host["pose_off"], host["point_off"], host["edge_off"], host["poses_init"], host["points_init"],
                           host["edge_pose"], host["edge_point"], host["edge_uv"])
for it in range(4):
    t0 = time.perf_counter(); setp(); t1 = time.perf_counter(); ba.upload(); torch.cuda.synchronize(); t2 = time.perf_counter()
    s = ba.optimize_resident(10); t3 = time.perf_counter()
    setp(); t4 = time.perf_counter(); s2 = ba.Optimize(10); t5 = time.perf_counter()
    print("set_problem %.2f ms | upload %.2f ms | optimize_resident %.2f ms (device %.2f) || set_problem %.2f | Optimize (upload+LM+readback) %.2f ms (device %.2f, setup %.2f)"
          % (1e3*(t1-t0), 1e3*(t2-t1), 1e3*(t3-t2), s.device_ms, 1e3*(t4-t3), 1e3*(t5-t4), s2.device_ms, s2.setup_ms))
