#!/usr/bin/env python
"""Parity of the SHARDED path (one process per GPU, points sharded, reduced camera system combined over
NCCL) — test infrastructure shared by tests/test_gpu_sharded.py, bench.py (world > 1) and tools/shard_check.py.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tests/shard_parity.py [--config 3|4] [--points 20000] [--poses 60]

Every rank holds the same global problem; rank 0 compares the refined state with (a) the CPU oracle, (b) the
same problem solved on one GPU without sharding, and all ranks check that they hold the same bits.  Prints one
JSON line on rank 0; exit code 1 on mismatch.  The reference concept being sharded is the Schur-reduced solve of
g2o's BlockSolver / LinearSolverDense (reference include/bundle_adjuster.h:108-117)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

COST_RTOL, POSE_ATOL, POINT_ATOL = 1e-5, 1e-6, 1e-5  # BASELINE.json north_star


def check(vb, torch, dist, rank, world, local, config=3, poses=60, points=20000, iters=6, huber=None, oracle_threads=8):
    """Collective: every rank of the process group calls it.  Returns the result dict on rank 0, None elsewhere."""
    p = vb.synth_config(config, n_poses=poses, n_points=points, target_edges=10 * points)
    hub = p["huber_delta"] if huber is None else huber
    a = (p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
    ba = vb.BundleAdjuster(len(a[1]), len(a[0]), len(a[2]), huber_delta=hub, device=local)
    ident = [vb.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ident, src=0)
    ba.shard_init(rank, world, ident[0])
    ba.set_problem(*a)
    s = ba.Optimize(iters)
    got_poses, got_points = ba.poses(), ba.points()
    # every rank must hold the same refined state, bit for bit
    t = torch.from_numpy(np.concatenate([got_poses.ravel(), got_points.ravel()])).cuda()
    ref = t.clone()
    dist.broadcast(ref, src=0)
    flags = torch.tensor([1.0 if bool(torch.equal(t, ref)) else 0.0], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    ba.close()
    out = None
    if rank == 0:
        import oracle_lib as orc
        c = orc.cfg(K=p["K"], huber_delta=hub, threads=oracle_threads)
        o_poses, o_points, fc, it, _ = orc.optimize(c, *a, iters)
        one = vb.BundleAdjuster(len(a[1]), len(a[0]), len(a[2]), huber_delta=hub, device=local)
        one.set_problem(*a)
        s1 = one.Optimize(iters)
        sign = np.sign(np.sum(got_poses[:, :4] * o_poses[:, :4], axis=1))[:, None]
        out = {
            "world": world, "layout": "config %d" % config, "poses": poses, "points": points, "observations": len(a[2]),
            "reduced_system_n": 6 * (poses - 1), "iterations": s.iterations, "oracle_iterations": it,
            "final_cost": s.final_cost, "oracle_cost": fc, "one_gpu_cost": s1.final_cost,
            "cost_rel_vs_oracle": abs(s.final_cost - fc) / fc,
            "pose_abs_vs_oracle": float(max(np.abs(got_poses[:, :4] * sign - o_poses[:, :4]).max(),
                                            np.abs(got_poses[:, 4:] - o_poses[:, 4:]).max())),
            "point_abs_vs_oracle": float(np.abs(got_points - o_points).max()),
            "pose_abs_vs_one_gpu": float(np.abs(got_poses - one.poses()).max()),
            "point_abs_vs_one_gpu": float(np.abs(got_points - one.points()).max()),
            "ranks_bit_identical": bool(flags.item() == 1.0),
            "device_ms": s.device_ms, "one_gpu_device_ms": s1.device_ms,
            "tolerances": {"cost_rel": COST_RTOL, "pose_abs": POSE_ATOL, "point_abs": POINT_ATOL},
        }
        out["ok"] = bool(out["cost_rel_vs_oracle"] < COST_RTOL and out["pose_abs_vs_oracle"] < POSE_ATOL
                         and out["point_abs_vs_oracle"] < POINT_ATOL and out["ranks_bit_identical"] and s.iterations == it)
        one.close()
    dist.barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=20000)
    ap.add_argument("--poses", type=int, default=60)
    ap.add_argument("--iters", type=int, default=6)
    ap.add_argument("--config", type=int, default=3, help="synthetic layout: 3 = grid (dense S), 4 = trajectory (banded S)")
    ap.add_argument("--out", default=None, help="also write the JSON line to this file (rank 0)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import viso_b200 as vb

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = check(vb, torch, dist, rank, world, local, args.config, args.poses, args.points, args.iters)
    ok = True
    if rank == 0:
        ok = out["ok"]
        line = json.dumps(out)
        print(line, flush=True)
        if args.out:
            with open(args.out, "w") as fh:
                fh.write(line + "\n")
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
