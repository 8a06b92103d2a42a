"""Would the small kernels of an LM trial (finalize, solve, decision: latency-bound, 117 us of 1504) hide behind the tile kernels
if the batch ran as two halves on two streams?  1024 C5 windows on one handle against 2 x 512 on two handles driven by two
host threads, their streams at equal and at different priorities (viso_ba_set_stream takes the caller's stream)."""
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import viso_b200 as vb

N = 1024
probs = [vb.synth_config(5, w) for w in range(N)]


def handle(ps, stream=None):
    b = vb.batch_problems(ps)
    ba = vb.BundleAdjuster(len(b["points_init"]), len(b["poses_init"]), len(b["edge_pose"]), device=0)
    if stream is not None:
        ba.set_stream(stream.cuda_stream)
    ba.set_problem_batched(b["pose_off"], b["point_off"], b["edge_off"], b["poses_init"], b["points_init"], b["edge_pose"], b["edge_point"], b["edge_uv"])
    ba.upload()
    return ba


def timed(handles, reps=7):
    walls = []
    for _ in range(reps):
        for h in handles:
            h.reset_state()
        torch.cuda.synchronize()
        bar = threading.Barrier(len(handles) + 1)
        out = [None] * len(handles)

        def work(i):
            bar.wait()
            out[i] = handles[i].optimize_resident(10)

        th = [threading.Thread(target=work, args=(i,)) for i in range(len(handles))]
        for t in th:
            t.start()
        bar.wait()
        t0 = time.perf_counter()
        for t in th:
            t.join()
        walls.append(1e3 * (time.perf_counter() - t0))
    return float(np.median(walls)), [o.device_ms for o in out], sum(o.trials for o in out) if out[0] is not None else 0


torch.cuda.set_device(0)
full = handle(probs)
w, dev, tr = timed([full])
print("one handle, 1024 windows:           wall %.2f ms, device %s" % (w, ["%.2f" % d for d in dev]), flush=True)
full.close()
for name, pa, pb in (("equal priority", 0, 0), ("first half high priority", -1, 0)):
    sa, sb = torch.cuda.Stream(priority=pa), torch.cuda.Stream(priority=pb)
    A, B = handle(probs[: N // 2], sa), handle(probs[N // 2:], sb)
    w, dev, tr = timed([A, B])
    print("two handles of 512, %-25s wall %.2f ms, device %s" % (name + ":", w, ["%.2f" % d for d in dev]), flush=True)
    A.close(), B.close()
