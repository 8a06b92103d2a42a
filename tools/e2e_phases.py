"""Where the end-to-end step of the C5 workload spends its wall time: set_problem (H2D + structure build) against
Optimize (LM loop + read-back), one handle and two handles on two threads, more steps than bench.py's default so
that the pipeline's fill and drain do not dominate.  GPU box only; prints one JSON object."""
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
import viso_b200 as vb  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 12
    probs, huber = bench.make_workload(vb, "c5", 0, 1024)
    batch = vb.batch_problems(probs)
    keep, host = [], {}
    for k, v in batch.items():
        t, a = bench.pinned_copy(torch, v)
        keep.append(t)
        host[k] = a
    out = {"steps": steps}
    ba = bench.make_adjuster(vb, host, huber, 0)
    for _ in range(2):
        bench.set_problem(ba, host)
        ba.Optimize(10)
    t_set = t_opt = 0.0
    dev = setup = 0.0
    for _ in range(steps):
        t0 = time.perf_counter()
        bench.set_problem(ba, host)
        t1 = time.perf_counter()
        s = ba.Optimize(10)
        t2 = time.perf_counter()
        t_set += t1 - t0
        t_opt += t2 - t1
        dev += s.device_ms
        setup += s.setup_ms
    out["one_handle"] = {"set_problem_ms": 1e3 * t_set / steps, "optimize_ms": 1e3 * t_opt / steps,
                         "device_lm_ms": dev / steps, "setup_ms_inside_optimize": setup / steps}
    ba2 = bench.make_adjuster(vb, host, huber, 0)
    hs = (ba, ba2)
    acc = [[0.0, 0.0], [0.0, 0.0]]
    trace, T0 = [], [time.perf_counter()]

    def worker(j, n):
        for _ in range(n):
            t0 = time.perf_counter()
            bench.set_problem(hs[j], host)
            t1 = time.perf_counter()
            sm = hs[j].Optimize(10)
            t2 = time.perf_counter()
            acc[j][0] += t1 - t0
            acc[j][1] += t2 - t1
            trace.append((j, round(1e3 * (t0 - T0[0]), 2), round(1e3 * (t1 - T0[0]), 2), round(1e3 * (t2 - T0[0]), 2),
                          round(sm.setup_ms, 2), round(sm.device_ms, 2)))
    for j in (0, 1):
        worker(j, 1)
    acc[0][:] = acc[1][:] = [0.0, 0.0]
    del trace[:]
    T0[0] = t0 = time.perf_counter()
    th = [threading.Thread(target=worker, args=(j, steps // 2)) for j in (0, 1)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    w = time.perf_counter() - t0
    n = 2 * (steps // 2)
    out["two_handles"] = {"ms_per_step": 1e3 * w / n, "set_problem_ms": 1e3 * (acc[0][0] + acc[1][0]) / n,
                          "optimize_ms": 1e3 * (acc[0][1] + acc[1][1]) / n,
                          "trace_handle_set0_set1_opt1_setupms_devicems": sorted(trace, key=lambda r: r[1])}
    # the LM loops take turns (a lock around the GPU-bound call), so that the two threads cannot fall into step
    for name, early_upload in (("two_handles_lm_lock", False), ("two_handles_lm_lock_upload_outside", True)):
        lock = threading.Lock()

        def worker2(j, n):
            for _ in range(n):
                bench.set_problem(hs[j], host)
                if early_upload:
                    hs[j].upload()
                with lock:
                    hs[j].Optimize(10)
        for j in (0, 1):
            worker2(j, 1)
        t0 = time.perf_counter()
        th = [threading.Thread(target=worker2, args=(j, steps // 2)) for j in (0, 1)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        out[name] = {"ms_per_step": 1e3 * (time.perf_counter() - t0) / n}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
