#!/usr/bin/env python
"""bench.py — bundle-adjustment throughput on B200 (BASELINE.json metric: BA observations/sec
and LM iterations/sec), one JSON line on stdout.

    python bench.py --gpus N --steps K --warmup W [--workload c5|c1|c2|c3|c4] [--impl reference]

A "step" is one bundle adjustment of the workload: `--lm-iters` (10) Levenberg-Marquardt
iterations from the same perturbed start.  observations/sec = sum over problems of
(edges x LM trials executed) / time, a trial being one linearise + Schur + solve + update +
cost evaluation.

  value : device-resident leg (inputs already in HBM), CUDA-event time of the LM loop on the
          stream the kernels are launched on, max over ranks.
  e2e   : the same metric through the C ABI with HOST buffers: set_problem (pinned host arrays)
          + Optimize (H2D, structure build, LM, D2H of poses and points), wall clock; timed on one
          handle and as a stream of steps through two handles (viso_b200.optimize_stream), the
          faster one reported if both give the identical final cost (`e2e.mode`).
  --impl reference : the CPU oracle (the reference adjuster itself cannot be compiled, see
          DESIGN.md) on all host cores over a bounded sample of the same workload.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "ba_observations_per_sec"
UNIT = "obs/s"
WORKLOADS = {
    "c1": "BASELINE configs[0]: one window, 10 keyframes / 2k points / ~15k observations, L2",
    "c2": "BASELINE configs[1]: one sliding window, 20 keyframes / 10k points / ~80k observations, Huber",
    "c3": "BASELINE configs[2]: global BA, 200 keyframes / 100k points / ~1M observations (sharded by point for N>1)",
    "c4": "BASELINE configs[3]: large-scale BA, 2k cameras / 1M points / ~10M observations (sharded by point for N>1)",
    "c5": "BASELINE configs[4]: 1024 concurrent 10-keyframe / 2k-point / ~15k-observation windows per GPU",
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c5", choices=sorted(WORKLOADS))
    ap.add_argument("--lm-iters", type=int, default=10)
    ap.add_argument("--windows", type=int, default=1024, help="windows per GPU (workload c5)")
    ap.add_argument("--cpu-sample-s", type=float, default=12.0, help="target seconds of the CPU baseline sample")
    ap.add_argument("--no-others", action="store_true", help="skip the secondary workloads in the JSON line")
    ap.add_argument("--sharded", default="c3,c4", help="N > 1: workloads run point-sharded beside the headline line ('' = none)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (profiling runs)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (profiling runs)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------------------------
def make_workload(vb, name, rank, windows):
    """returns (list of problems, huber delta, sharded?)"""
    if name == "c5":
        probs = [vb.synth_config(5, rank * windows + w) for w in range(windows)]
        return probs, 0.0
    cid = {"c1": 1, "c2": 2, "c3": 3, "c4": 4}[name]
    if name in ("c3", "c4"):
        p = vb.synth_config(cid)  # every rank holds the SAME problem; it is sharded by point
    else:
        sp = vb.synth_preset(cid)
        sp.seed += 0x1000 * rank  # replicas: every GPU its own window
        p = vb.synth_problem(sp)
    return [p], p["huber_delta"]


def pinned_copy(torch, a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t, t.numpy()


def pinned_batch(torch, batch):
    """the arrays of a batch in page-locked host memory (what a caller that cares about H2D time hands over)"""
    keep, host = [], {}
    for k, v in batch.items():
        t, a = pinned_copy(torch, v)
        keep.append(t)
        host[k] = a
    host["_keep"] = keep
    return host


def make_adjuster(vb, batch, huber, device):
    ba = vb.BundleAdjuster(len(batch["points_init"]), len(batch["poses_init"]), len(batch["edge_pose"]),
                           huber_delta=huber, device=device)
    return ba


def set_problem(ba, batch):
    if len(batch["pose_off"]) > 2:
        ba.set_problem_batched(batch["pose_off"], batch["point_off"], batch["edge_off"], batch["poses_init"],
                               batch["points_init"], batch["edge_pose"], batch["edge_point"], batch["edge_uv"])
    else:
        ba.set_problem(batch["poses_init"], batch["points_init"], batch["edge_pose"], batch["edge_point"], batch["edge_uv"])


def units_of(ba, batch):
    """sum over problems of edges x trials of the last optimize call"""
    sums = ba.problem_summaries()
    ne = np.diff(batch["edge_off"])
    return float(sum(int(n) * s.trials for n, s in zip(ne, sums))), sum(s.trials for s in sums)


# ---------------------------------------------------------------------------------------------
# clocks during the timed region (pynvml; falls back to nvidia-smi)
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.stop_flag, self.max_mhz = [], set(), False, None
        self.thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for b, n in self.REASONS.items():
                    if bits & b and n != "gpu_idle":
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        if self.nv:
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------
# CPU baseline: the oracle on all host cores, one problem per thread (windows are independent)
# ---------------------------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_leg(vb, workload, lm_iters, target_s, windows):
    """returns dict(value, unit, cores, kind, sample, seconds, units)"""
    import oracle_lib as orc
    from concurrent.futures import ThreadPoolExecutor
    cores = host_cores()

    def solve(p, threads):
        c = orc.cfg(K=p["K"], huber_delta=p["huber_delta"], threads=threads)
        _, _, _, _, st = orc.optimize(c, p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"], lm_iters)
        return len(p["edge_pose"]) * sum(s.trials for s in st)

    single = None
    if workload == "c5":
        probe = vb.synth_config(5, 0)
        t0 = time.perf_counter()
        u1 = solve(probe, 1)
        t1 = time.perf_counter() - t0
        # the reference and default g2o are single-threaded: one window on one thread, beside the all-cores figure
        single = {"value": u1 / t1, "unit": UNIT, "cores": 1, "kind": "port",
                  "sample": "window 0 of the workload on one thread, %d LM iterations" % lm_iters}
        n = int(max(cores, min(windows, cores * max(1.0, target_s / max(t1, 1e-3)))))
        n = (n // cores) * cores
        probs = [vb.synth_config(5, w) for w in range(n)]
        t0 = time.perf_counter()
        with ThreadPoolExecutor(cores) as ex:  # ctypes releases the GIL: `cores` windows in flight
            units = sum(ex.map(lambda q: solve(q, 1), probs))
        dt = time.perf_counter() - t0
        sample = "%d of the %d windows, one window per thread, %d LM iterations each" % (n, windows, lm_iters)
    else:
        if workload == "c4":
            raise SystemExit("the CPU oracle factorises the reduced system densely (n = 11994: minutes per LM iteration); "
                             "no bounded CPU sample of workload c4 exists, use c3 or c5")
        p = vb.synth_config({"c1": 1, "c2": 2, "c3": 3}[workload])
        t0 = time.perf_counter()
        u1 = solve(p, 1)
        t1 = time.perf_counter() - t0
        single = {"value": u1 / t1, "unit": UNIT, "cores": 1, "kind": "port",
                  "sample": "the whole problem once on one thread, %d LM iterations" % lm_iters}
        reps, units, dt = 0, 0.0, 0.0
        while dt < target_s and reps < 200:
            t0 = time.perf_counter()
            units += solve(p, cores)
            dt += time.perf_counter() - t0
            reps += 1
        sample = "the whole problem, %d repetitions of %d LM iterations, OpenMP over points" % (reps, lm_iters)
    return {"value": units / dt, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
            "seconds": round(dt, 3), "units": units, "single_thread": single}


def run_reference(args):
    """--impl reference: the CPU oracle port is the only runnable statement of the reference path
    (the reference adjuster does not compile; no oracle/_ref exists).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import viso_b200 as vb
    per = max(2.0, min(args.cpu_sample_s, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_leg(vb, args.workload, args.lm_iters, per / 4, args.windows)
    units, secs, last = 0.0, 0.0, None
    for _ in range(args.steps):
        last = cpu_leg(vb, args.workload, args.lm_iters, per, args.windows)
        units += last["units"]
        secs += last["seconds"]
    value = units / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "description": WORKLOADS[args.workload], "lm_iterations": args.lm_iters,
                   "note": "each step is a bounded sample of the workload on the host cores"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": last["cores"], "kind": "port", "sample": last["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------
# roofline of the dominant kernel: algorithmic bytes per launch (DESIGN.md "Algorithmic bytes")
# ---------------------------------------------------------------------------------------------
def algorithmic_bytes(No, Np, Nc, Nblocks):
    """SURVEY 8(d) per LM iteration, split by pass.  Storage: observation = 16 B (u, v fp32 + two
    int32 indices), point = 24 B, per-point saved (Hll^-1, g_l) = 72 B, S block = 288 B, pose = 96 B."""
    S = 288.0 * Nblocks + 48.0 * Nc
    return {
        "L": 16.0 * No + (24.0 + 72.0) * Np + 96.0 * Nc + S,           # linearise + Schur
        "B": 16.0 * No + (24.0 + 72.0 + 24.0) * Np + 48.0 * Nc,        # back-substitute + update
        "C": 16.0 * No + 24.0 * Np + 96.0 * Nc,                         # cost at the candidate
    }


# which pass each kernel belongs to (kernels of one pass are timed together)
PASS_OF = {"linearize": "L", "schur_blocks": "L", "block_finalize": "L", "win_linearize_schur": "L", "win_finalize": "L",
           "update": "BC", "win_update": "BC", "solve_small": None, "solve_large": None, "reduce": None, "decide": None}


# FP64 instructions (DFMA / DMUL / DADD ..., one issue slot each) per observation of one launch, derived by
# tools/ncu_derive.py from an `ncu --set full` capture of THIS build (profiles/derived_roofline.json: keyed by the
# source digest in viso_b200/.libviso_ba.stamp; `stale` is reported when the library has changed since the capture).
# The path is bound by the FP64 pipe, not by HBM, so this second roof is reported beside the first.
def derived_profile():
    try:
        with open(os.path.join(ROOT, "profiles", "derived_roofline.json")) as fh:
            d = json.load(fh)
    except Exception:
        return None
    try:
        with open(os.path.join(ROOT, "viso_b200", ".libviso_ba.stamp")) as fh:
            d["stale"] = fh.read().strip() != d.get("source_digest")
    except Exception:
        d["stale"] = None
    return d


def fp64_peak():
    """measured FP64 FMA issue rate (profiles/r1_fma_peaks.json, tools/peaks.cu): T slots/s"""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_fma_peaks.json")) as fh:
            return float(json.load(fh)["fp64_fma_tflops"]) / 2.0, "measured (profiles/r1_fma_peaks.json)"
    except Exception:
        return 148 * 64 * 1.965e9 / 1e12, "nominal 64 DFMA/clk/SM x 148 SMs x 1.965 GHz"


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def roofline(ktimes, trials_per_problem, sizes, nprob=1):
    """ktimes: name -> (launches, total ms) of ONE optimize call timed kernel by kernel."""
    No, Np, Nc, Nblocks = sizes
    A = algorithmic_bytes(No, Np, Nc, Nblocks)
    A["BC"] = A["B"] + A["C"]
    A_iter = A["L"] + A["B"] + A["C"]
    peak, peak_src = measured_peak()
    per_kernel = {k: {"launches": n, "avg_us": 1e3 * ms / max(1, n), "total_ms": ms} for k, (n, ms) in ktimes.items()}
    top = max(ktimes, key=lambda k: ktimes[k][1])
    ps = PASS_OF.get(top)
    if ps is None:  # a solve kernel dominates: it has no HBM-bound share worth a roofline; use pass L anyway
        ps = "L"
    members = [k for k in ktimes if PASS_OF.get(k) == ps]
    # Units one launch processes: a launch of the pass over EVERY problem moves A[ps] algorithmic bytes.  In a
    # batch, the straggler launches after the first batch of trials cover only the few problems still active,
    # so the pass's time is divided by the number of full-launch equivalents it actually processed
    # (problem-trials / problems, plus the init linearisation for pass L) -- not by the launch count, which
    # would credit the near-empty straggler launches with a full launch's bytes.  One problem: the same thing.
    n_equiv = trials_per_problem + (1.0 if ps == "L" else 0.0)
    t_us = 1e3 * sum(ktimes[k][1] for k in members) / max(n_equiv, 1e-9)
    achieved = A[ps] / (t_us * 1e-6) / 1e9
    # measured DRAM traffic of the same kernels (ncu --set full, one launch over `windows` windows of the
    # same shape): scaled to this run's window count, per launch like `achieved`
    traffic, fp64 = None, None
    prof = derived_profile()
    pk_ = (prof or {}).get("kernels", {})
    if prof and all(k in pk_ for k in members) and prof.get("problems"):
        # dram__bytes_read + dram__bytes_write of the pass's kernels in the captured launch, scaled to this run's
        # problem count (the capture is of the default bench command, so the scale is normally 1)
        traffic = sum(pk_[k]["dram_bytes"] for k in members) * nprob / float(prof["problems"])
    if prof and top in pk_ and pk_[top].get("fp64_inst_per_observation"):
        slots = float(pk_[top]["fp64_inst_per_observation"])
        pk, src = fp64_peak()
        n_top = trials_per_problem + (1.0 if PASS_OF.get(top) == "L" else 0.0)  # the init launch counted as a full one
        ach = slots * No * n_top / (ktimes[top][1] * 1e-3) / 1e12
        fp64 = {"kernel": top, "slots_per_observation": slots, "achieved": ach, "peak": pk,
                "unit": "T FP64 instructions/s (an FMA is one)", "frac": ach / pk, "peak_source": src,
                "capture": prof.get("capture"), "capture_is_of_this_build": (not prof["stale"]) if prof.get("stale") is not None else None,
                "note": "the binding roof of this kernel: the pair products are FP64-issue-bound, HBM is far from saturated"}
    return {
        "fp64": fp64,
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
        "kernel": top, "pass": ps, "pass_kernels": members, "pass_us_per_launch": t_us,
        "full_launch_equivalents": n_equiv,
        "algorithmic_bytes_per_launch": A[ps], "algorithmic_bytes_per_lm_iteration": A_iter,
        "peak_source": peak_src, "kernels": per_kernel,
    }


# ---------------------------------------------------------------------------------------------
# the B200 arm
# ---------------------------------------------------------------------------------------------
def bind_to_gpu_numa_node(torch, local):
    """N ranks on one box: keep this rank's host threads, and with them the page-locked buffers it allocates from here on
    (first touch), on the NUMA node its GPU hangs off.  Eight ranks x two host threads x 345 MB per step otherwise meet on
    whatever node the launcher left them on.  Best effort: returns the node, or None when sysfs does not say."""
    try:
        bus = torch.cuda.get_device_properties(local).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(local), "pci_domain_id", 0)
        dev = getattr(torch.cuda.get_device_properties(local), "pci_device_id", 0)
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    import viso_b200 as vb

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run --nproc-per-node %d"
                         % (args.gpus, world, args.gpus))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: no CUDA device is visible and this backend has no CPU fallback")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(torch, local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    sharded = args.workload in ("c3", "c4") and world > 1
    probs, huber = make_workload(vb, args.workload, rank, args.windows)
    batch = vb.batch_problems(probs)
    # pinned host copies: what the e2e leg hands to the C ABI every step
    keep, host = [], {}
    for k, v in batch.items():
        t, a = pinned_copy(torch, v)
        keep.append(t)
        host[k] = a
    No, Np, Nc = len(host["edge_pose"]), len(host["points_init"]), len(host["poses_init"])
    nprob = len(probs)
    nfree = [len(p["poses_init"]) - 1 for p in probs]
    Nblocks = sum(f * (f + 1) // 2 for f in nfree)

    ba = make_adjuster(vb, host, huber, local)
    if sharded:
        ident = [vb.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        ba.shard_init(rank, world, ident[0])
    set_problem(ba, host)
    ba.upload()

    flush = None
    working_set = 12.0 * No + 96.0 * No  # observations + per-observation records
    if working_set < 2 * 126e6:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def step_resident():
        if flush is not None:
            flush.zero_()
            torch.cuda.synchronize()
        ba.reset_state()
        return ba.optimize_resident(args.lm_iters)

    # ---- device-resident leg ----
    for _ in range(max(3, args.warmup)):
        s = step_resident()
    units_step, trials_step = units_of(ba, host)
    if sharded:
        units_job = units_step           # one problem, every rank reports the same trials
    else:
        units_job = sum_over_ranks(units_step)
    clocks = ClockSampler(local)
    launches0 = ba.kernel_launches()
    barrier()
    clocks.start()
    t_wall0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(args.steps):
        s = step_resident()
        dev_ms += s.device_ms
    barrier()
    wall_s = time.perf_counter() - t_wall0
    clk = clocks.stop()
    launches = ba.kernel_launches() - launches0
    dev_ms = max_over_ranks(dev_ms)
    ms_per_step = dev_ms / args.steps
    value = units_job / (ms_per_step * 1e-3)
    lm_trials_job = trials_step if sharded else sum_over_ranks(float(trials_step))

    # ---- kernel-by-kernel pass for the roofline (not part of any reported throughput) ----
    ba.reset_state()
    ba.set_kernel_timing(True)
    ba.optimize_resident(args.lm_iters)
    ktimes = ba.kernel_times()
    ba.set_kernel_timing(False)
    ktimes = {k: v for k, v in ktimes.items() if k not in ("state_init", "cam_rt", "lm_init", "gather_current")}
    roof = roofline(ktimes, trials_step / nprob, (No, Np, Nc, Nblocks), nprob)
    A_iter = roof["algorithmic_bytes_per_lm_iteration"]
    roof["iteration"] = {"achieved": A_iter * (trials_step / nprob) / (ms_per_step * 1e-3) / 1e9,
                         "frac": A_iter * (trials_step / nprob) / (ms_per_step * 1e-3) / 1e9 / roof["peak"],
                         "note": "whole LM loop incl. the reduced-system solve: A_iter x trials / device time"}

    # ---- end-to-end leg: host buffers in, refined poses / points out, every step ----
    h2d = sum(host[k].nbytes for k in ("poses_init", "points_init", "edge_pose", "edge_point", "edge_uv"))
    d2h = host["poses_init"].nbytes + host["points_init"].nbytes
    e2e_steps = 0 if args.no_e2e else args.steps
    for _ in range(0 if args.no_e2e else 2):
        set_problem(ba, host)
        ba.Optimize(args.lm_iters)
    barrier()
    t0 = time.perf_counter()
    loss = None
    for _ in range(e2e_steps):
        set_problem(ba, host)
        s_e2e = ba.Optimize(args.lm_iters)
        loss = s_e2e.final_cost  # the step's result, read on the host
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_serial_ms = 1e3 * e2e_s / max(1, e2e_steps)
    e2e_mode = "one handle, steps back to back"
    if e2e_steps and not sharded:
        # The same steps as a stream of batches through TWO handles on two host threads (viso_b200/stream.py:
        # ctypes drops the GIL in the C ABI; one handle = one stream, include/viso_ba.h): a step's H2D / structure
        # build / D2H overlap the other handle's LM loop, the LM loops take turns.  Every step still uploads its
        # own inputs and reads its own result.
        try:
            ba2 = make_adjuster(vb, host, huber, local)
            losses = []
            vb.optimize_stream((ba, ba2), [host, host], args.lm_iters)  # warm the second handle's arena
            barrier()
            t0 = time.perf_counter()
            vb.optimize_stream((ba, ba2), [host] * e2e_steps, args.lm_iters,
                               consume=lambda i, h, s: losses.append(s.final_cost))
            barrier()
            piped_s = max_over_ranks(time.perf_counter() - t0)
            if len(losses) == e2e_steps and all(x == loss for x in losses) and piped_s < e2e_s:
                e2e_s = piped_s
                e2e_mode = ("two handles on two host threads (viso_b200.optimize_stream): a step's transfers and "
                            "structure build overlap the other step's LM loop")
            ba2.close()
        except Exception as e:  # the one-handle number stands
            e2e_mode += " (two-handle pipeline failed: %s)" % e
    e2e_value = units_job * e2e_steps / e2e_s if e2e_steps else None
    loss = loss if e2e_steps else None

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if sharded else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {
            "workload": args.workload, "description": WORKLOADS[args.workload], "lm_iterations": args.lm_iters,
            "problems_per_gpu": nprob, "observations_per_gpu": No, "points_per_gpu": Np, "poses_per_gpu": Nc,
            "parallelism": ("point-sharded x%d + fp64 NCCL allreduce" % world) if sharded else ("replicas x%d" % world),
            "l2": "flushed between steps (256 MiB write)" if flush is not None else "working set larger than L2 (126 MB)",
            "host_numa_node": numa,
        },
        "lm_iters_per_sec": lm_trials_job / (ms_per_step * 1e-3),
        "lm_trials_per_step": lm_trials_job, "final_cost": s.final_cost,
        "wall_ms_per_step": 1e3 * wall_s / args.steps,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": 1e3 * e2e_s / max(1, e2e_steps), "final_cost": loss,
                "mode": e2e_mode, "ms_per_step_one_handle": e2e_serial_ms},
        "gpu_launches": int(launches), "clocks": clk, "roofline": roof,
    }
    if rank == 0 and not args.no_others and world == 1:
        line["others"] = others(vb, args, local)
    if world > 1 and args.sharded and not sharded:
        ba.close()  # the headline handle's arena goes back before the large problems are set up
        sh = sharded_section(vb, torch, dist, args, rank, world, local, barrier, max_over_ranks)
        if rank == 0:
            line["sharded"] = sh
    if rank == 0 and not args.no_cpu and args.workload != "c4":
        line["cpu_baseline"] = {k: v for k, v in cpu_leg(vb, args.workload, args.lm_iters, args.cpu_sample_s, args.windows).items()
                                if k in ("value", "unit", "cores", "kind", "sample", "single_thread")}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def sharded_section(vb, torch, dist, args, rank, world, local, barrier, max_over_ranks):
    """N > 1 only (collective: every rank calls it).  The workloads the path SHARDS -- BASELINE configs[2] (c3) and
    configs[3] (c4): one problem, points sharded over the ranks, the reduced camera system combined over NCCL on the
    handle's own communicator -- run on the process group of the headline line, so that the driver's scaling record
    carries a strong-scaling curve, beside the same problem solved by rank 0 alone on its GPU, and a parity check of the
    sharded path against the CPU oracle at a bounded size (tests/shard_parity.py).  Reference concept being sharded:
    BlockSolver<6,3> + LinearSolverDense, reference include/bundle_adjuster.h:108-117."""
    out = {}
    names = [w for w in ("c3", "c4") if w in args.sharded.split(",")]
    for name in names:
        try:
            probs, huber = make_workload(vb, name, rank, 0)
            b = pinned_batch(torch, vb.batch_problems(probs))
            No = len(b["edge_pose"])
            ba = make_adjuster(vb, b, huber, local)
            ident = [vb.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ident, src=0)
            ba.shard_init(rank, world, ident[0])
            t0 = time.perf_counter()
            set_problem(ba, b)
            ba.upload()
            setup_s = max_over_ranks(time.perf_counter() - t0)
            for _ in range(3):
                ba.reset_state()
                s = ba.optimize_resident(args.lm_iters)
            steps = max(3, min(args.steps, 10 if name == "c3" else 5))
            barrier()
            ms = 0.0
            for _ in range(steps):
                ba.reset_state()
                ms += ba.optimize_resident(args.lm_iters).device_ms
            barrier()
            ms = max_over_ranks(ms) / steps
            ba.reset_state()
            ba.set_kernel_timing(True)
            ba.optimize_resident(args.lm_iters)
            kt = ba.kernel_times()
            ba.set_kernel_timing(False)
            # end to end through the C ABI with host buffers: every rank hands over the problem, reads the result back
            # (page-locked arrays, as in the headline's e2e leg; the median of three calls, each the max over the ranks)
            walls = []
            for _ in range(3):
                barrier()
                t0 = time.perf_counter()
                set_problem(ba, b)
                s_e = ba.Optimize(args.lm_iters)
                barrier()
                walls.append(max_over_ranks(time.perf_counter() - t0))
            e2e_s = float(np.median(walls))
            # every rank must hold the same refined state, bit for bit
            t = torch.from_numpy(np.concatenate([ba.poses().ravel(), ba.points().ravel()])).cuda()
            ref = t.clone()
            dist.broadcast(ref, src=0)
            same = torch.tensor([1.0 if bool(torch.equal(t, ref)) else 0.0], device="cuda")
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            ba.close()
            rec = {"value": No * s.trials / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "lm_trials_per_step": s.trials,
                   "lm_iters_per_sec": s.trials / (ms * 1e-3), "observations": No, "steps": steps, "scaling": "strong",
                   "final_cost": s.final_cost, "ranks_bit_identical": bool(same.item() == 1.0), "setup_s": setup_s,
                   "e2e": {"value": No * s_e.trials / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s, "final_cost": s_e.final_cost},
                   "kernels_ms_per_step": {k: v[1] for k, v in kt.items() if k not in ("state_init", "cam_rt", "lm_init", "gather_current")}}
            if rank == 0:  # the same problem on ONE GPU, measured in the same process on the same box
                one = make_adjuster(vb, b, huber, local)
                set_problem(one, b)
                one.upload()
                for _ in range(2):
                    one.reset_state()
                    s1 = one.optimize_resident(args.lm_iters)
                m1 = []
                for _ in range(3):
                    one.reset_state()
                    m1.append(one.optimize_resident(args.lm_iters).device_ms)
                one.close()
                m1 = float(np.median(m1))
                rec["one_gpu"] = {"ms_per_step": m1, "value": No * s1.trials / (m1 * 1e-3), "final_cost": s1.final_cost,
                                  "cost_rel_vs_sharded": abs(s1.final_cost - s.final_cost) / s1.final_cost}
                rec["speedup_vs_one_gpu"] = m1 / ms
            barrier()
            out[name] = rec
        except Exception as e:  # never fail the headline because of it
            out[name] = {"error": str(e)}
            try:
                barrier()
            except Exception:
                pass
    try:
        import shard_parity
        out["parity"] = {
            "dense_S": shard_parity.check(vb, torch, dist, rank, world, local, config=3, poses=60, points=20000, iters=6),
            "banded_S": shard_parity.check(vb, torch, dist, rank, world, local, config=4, poses=150, points=15000, iters=5),
        }
    except Exception as e:
        out["parity"] = {"error": str(e)}
    return out


def others(vb, args, device):
    """device-resident numbers of the other single-GPU configurations (parity-test cases; context only)"""
    import torch
    out = {}
    for name in ("c1", "c2", "c3", "c4"):
        if name == args.workload:
            continue
        try:
            probs, huber = make_workload(vb, name, 0, 0)
            b = pinned_batch(torch, vb.batch_problems(probs))
            ba = make_adjuster(vb, b, huber, device)
            set_problem(ba, b)
            ba.upload()
            for _ in range(3):
                ba.reset_state()
                s = ba.optimize_resident(args.lm_iters)
            ms = []
            for _ in range(5):
                ba.reset_state()
                ms.append(ba.optimize_resident(args.lm_iters).device_ms)
            m = float(np.median(ms))
            out[name] = {"value": len(b["edge_pose"]) * s.trials / (m * 1e-3), "unit": UNIT, "ms_per_step": m,
                         "us_per_lm_iteration": 1e3 * m / max(1, s.trials),
                         "lm_iters_per_sec": s.trials / (m * 1e-3), "observations": len(b["edge_pose"]),
                         "final_cost": s.final_cost, "note": "device-resident, L2 not flushed"}
            if name in ("c1", "c2"):  # the VO call (reference src/viso.cpp:121): one window per Optimize, host buffers in and out
                wall = []
                for _ in range(5):
                    t0 = time.perf_counter()
                    set_problem(ba, b)
                    se = ba.Optimize(args.lm_iters)
                    wall.append(time.perf_counter() - t0)
                w = float(np.median(wall))
                out[name]["e2e"] = {"value": len(b["edge_pose"]) * se.trials / w, "unit": UNIT, "ms_per_step": 1e3 * w,
                                    "final_cost": se.final_cost, "note": "set_problem + Optimize with host arrays: H2D, structure build, LM, D2H"}
            ba.close()
        except Exception as e:  # context only: never fail the headline because of it
            out[name] = {"error": str(e)}
    try:
        out["photo_ba"] = photo_leg(vb, device, args.lm_iters, cpu=not args.no_cpu)
    except Exception as e:
        out["photo_ba"] = {"error": str(e)}
    try:
        out["direct_pose"] = direct_leg(device, cpu=not args.no_cpu)
    except Exception as e:
        out["direct_pose"] = {"error": str(e)}
    try:
        out["lk_align"] = lk_leg(device, cpu=not args.no_cpu)
    except Exception as e:
        out["lk_align"] = {"error": str(e)}
    return out


def photo_leg(vb, device, lm_iters, cpu=True):
    """LM over the reference adjuster's photometric edge (reference include/bundle_adjuster.h:58-106): a
    window of 6 keyframes (320 x 240) and 1500 points, 16 samples per edge.  Unit: edges x LM trials per second."""
    from viso_b200.direct_synth import synth_photo_window
    win = synth_photo_window(seed=9, n_frames=6, n_points=1500)
    K = win["K"]
    ba = vb.BundleAdjuster(len(win["points_init"]), len(win["poses_init"]), len(win["edge_point"]), device=device,
                           fx=K[0], fy=K[1], cx=K[2], cy=K[3])
    ba.set_photo_problem(win["poses_init"], win["points_init"], win["edge_point"], win["edge_src"], win["edge_dst"], win["images"])
    ba.SetPoseFixed(0, True), ba.SetPoseFixed(1, True)
    ba.upload()
    ms = []
    for i in range(8):
        ba.reset_state()
        s = ba.optimize_resident(lm_iters)
        if i >= 3:
            ms.append(s.device_ms)
    m = float(np.median(ms))
    ne = len(win["edge_point"])
    res = {"edges": ne, "samples_per_edge": 16, "keyframes": len(win["poses_init"]), "points": len(win["points_init"]),
           "unit": "edges x LM trials / s", "value": ne * s.trials / (m * 1e-3), "ms_per_step": m, "lm_trials_per_step": s.trials,
           "initial_cost": s.initial_cost, "final_cost": s.final_cost}
    ba.close()
    if cpu:
        import oracle_lib as orc
        import photo_ba_oracle_lib as porc
        sub = 150  # the oracle forms ONE dense system over poses and points: a bounded sample of the points
        sel = win["edge_point"] < sub
        t0 = time.perf_counter()
        _, _, fc, it, st = porc.optimize(orc.cfg(K=K), win["images"], win["poses_init"], win["points_init"][:sub], win["edge_point"][sel],
                                         win["edge_src"][sel], win["edge_dst"][sel], 3, fixed=[1, 1, 0, 0, 0, 0])
        dt = time.perf_counter() - t0
        tr = sum(x.trials for x in st)
        res["cpu_oracle"] = {"value": int(sel.sum()) * tr / dt, "unit": res["unit"], "cores": 1, "kind": "port",
                             "sample": "%d of the %d points, 3 LM iterations, dense normal equations" % (sub, len(win["points_init"]))}
    return res


def lk_leg(device, cpu=True, n_points=4000):
    """Lucas-Kanade patch alignment, reference src/viso.cpp:845-925 (LKAlignmentSingle over
    4 levels) and :259-391 (optical flow), on a synthetic 640x480 frame pair.  Unit: keypoints/s through all four
    pyramid levels; kernel time by CUDA events inside the C ABI call; `e2e` adds the H2D/D2H of the keypoints."""
    from viso_b200 import lk
    from viso_b200.direct_synth import synth_scene, FX, FY, CX, CY
    sc = synth_scene(seed=12, n_points=n_points)
    P = sc["points"]
    uv_ref = np.stack([FX * P[:, 0] / P[:, 2] + CX, FY * P[:, 1] / P[:, 2] + CY], axis=1)
    a = lk.LKAligner(max_keypoints=n_points, slots=2, device=device)
    a.set_pyramid(0, sc["ref_pyr"])
    a.set_pyramid(1, sc["cur_pyr"])
    res = {"keypoints": n_points, "patch_pixels": 64, "image": "640x480, 4 levels", "unit": "keypoints/s"}

    def timed(fn, reps=20):
        ms, wall = [], []
        for i in range(reps + 3):
            t0 = time.perf_counter()
            out = fn()
            dt = time.perf_counter() - t0
            if i >= 3:
                ms.append(out[2])
                wall.append(dt)
        m, w = float(np.median(ms)), float(np.median(wall))
        return {"kernel_ms": m, "value": n_points / (m * 1e-3), "e2e_ms": 1e3 * w, "e2e_value": n_points / w,
                "success_rate": float(np.mean(out[1]))}
    res["lk_alignment"] = timed(lambda: a.LKAlignment(0, 1, uv_ref, uv_ref))
    kp1 = uv_ref.astype(np.float32)
    res["optical_flow_forward"] = timed(lambda: a.OpticalFlowMultiLevel(0, 1, kp1, None, inverse=False))
    res["optical_flow_inverse"] = timed(lambda: a.OpticalFlowMultiLevel(0, 1, kp1, None, inverse=True))
    a.close()
    if cpu:
        import lk_oracle_lib as lko
        t0 = time.perf_counter()
        lko.align(lko.cfg(), [sc["ref_pyr"]], 0, sc["cur_pyr"], uv_ref, uv_ref)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": n_points / dt, "unit": "keypoints/s", "cores": 1, "kind": "port",
                               "sample": "LKAlignment of %d pairs, oracle/lk_oracle.c single-threaded like the reference" % n_points}
    return res


def direct_leg(device, cpu=True, n_points=2000, batch=128):
    """the tracker's direct pose estimation, reference src/viso.cpp:661-766, on a
    synthetic 640x480 frame pair with 2000 map points.  Unit: Gauss-Newton iterations/s (one iteration =
    one pass over all points x 64 patch pixels + the 6x6 solve); kernel time by CUDA events inside the
    C ABI call.  `reference_loop` is the loop as the reference writes it (one step per level), `repaired`
    iterates to convergence; `batched` runs `batch` independent frames in one launch."""
    from viso_b200 import direct
    from viso_b200.direct_synth import synth_scene
    sc = synth_scene(seed=11, n_points=n_points)
    res = {"points": n_points, "patch_pixels": 64, "image": "640x480, 4 levels", "unit": "GN iterations/s"}

    def run(frames, ref_term, reps=20):
        est = direct.DirectPoseEstimator(max_points=n_points, frames=frames, device=device, reference_termination=ref_term)
        for f in range(frames):
            est.set_pyramid(0, sc["ref_pyr"], frame=f)
            est.set_pyramid(1, sc["cur_pyr"], frame=f)
        est.set_points(sc["points"])
        init = np.tile(sc["identity"], (frames, 1))
        ms = []
        for i in range(reps + 3):
            pose, st, m = est.DirectPoseEstimationMultiLayer(init)
            if i >= 3:
                ms.append(m)
        its = sum(s.iterations for fr in st for s in fr)
        m = float(np.median(ms))
        est.close()
        return {"frames": frames, "gn_iterations": its, "kernel_ms": m, "value": its / (m * 1e-3),
                "us_per_iteration_per_frame": 1e3 * m * frames / its, "frames_per_sec": frames / (m * 1e-3),
                "pose_error": float(np.abs(np.reshape(pose, (-1, 7))[0] - sc["true_pose"]).max())}
    res["reference_loop"] = run(1, 1)
    res["repaired"] = run(1, 0)
    res["batched"] = run(batch, 0, reps=5)
    if cpu:
        import direct_oracle_lib as dorc
        t0 = time.perf_counter()
        _, st = dorc.multi_layer(dorc.cfg(reference_termination=0), sc["ref_pyr"], sc["cur_pyr"], sc["ref_pose"], sc["identity"], sc["points"])
        dt = time.perf_counter() - t0
        its = sum(s.iterations for s in st)
        res["cpu_baseline"] = {"value": its / dt, "unit": "GN iterations/s", "cores": 1, "kind": "port",
                               "sample": "one frame, repaired loop, %d iterations, oracle/direct_oracle.c single-threaded like the reference" % its}
    return res


def emit(line):
    """the ONE JSON line goes to the real stdout; everything else any library prints while the
    bench runs (NCCL prints its version banner on stdout) has been sent to stderr by main()"""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
