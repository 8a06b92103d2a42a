#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun on N GPUs of one box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/shard_check.py [--points 20000]

Every rank holds the same global problem (config-3 shape, scaled by --points), the observations
are sharded by point range, the reduced camera system is combined with the fp64 NCCL allreduce.
Rank 0 compares the refined state with (a) the CPU oracle and (b) the same problem solved on one
GPU without sharding.  Prints one JSON line; exit code 1 on mismatch."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=20000)
    ap.add_argument("--poses", type=int, default=60)
    ap.add_argument("--iters", type=int, default=6)
    ap.add_argument("--config", type=int, default=3, help="synthetic layout: 3 = grid (dense-ish S), 4 = trajectory (banded S)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import viso_b200 as vb

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    p = vb.synth_config(args.config, n_poses=args.poses, n_points=args.points, target_edges=10 * args.points)
    a = (p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
    ba = vb.BundleAdjuster(len(a[1]), len(a[0]), len(a[2]), huber_delta=p["huber_delta"], device=local)
    ident = [vb.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ident, src=0)
    ba.shard_init(rank, world, ident[0])
    ba.set_problem(*a)
    s = ba.Optimize(args.iters)
    poses, points = ba.poses(), ba.points()
    # every rank must hold the same refined state, bit for bit
    t = torch.from_numpy(np.concatenate([poses.ravel(), points.ravel()])).cuda()
    ref = t.clone()
    dist.broadcast(ref, src=0)
    same = bool(torch.equal(t, ref))
    flags = torch.tensor([1.0 if same else 0.0], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    ok, out = True, None
    if rank == 0:
        import oracle_lib as orc
        c = orc.cfg(K=p["K"], huber_delta=p["huber_delta"], threads=8)
        o_poses, o_points, fc, it, _ = orc.optimize(c, *a, args.iters)
        one = vb.BundleAdjuster(len(a[1]), len(a[0]), len(a[2]), huber_delta=p["huber_delta"], device=local)
        one.set_problem(*a)
        s1 = one.Optimize(args.iters)
        sign = np.sign(np.sum(poses[:, :4] * o_poses[:, :4], axis=1))[:, None]
        out = {
            "world": world, "observations": len(a[2]), "iterations": s.iterations, "oracle_iterations": it,
            "final_cost": s.final_cost, "oracle_cost": fc, "one_gpu_cost": s1.final_cost,
            "cost_rel_vs_oracle": abs(s.final_cost - fc) / fc,
            "pose_abs_vs_oracle": float(max(np.abs(poses[:, :4] * sign - o_poses[:, :4]).max(), np.abs(poses[:, 4:] - o_poses[:, 4:]).max())),
            "point_abs_vs_oracle": float(np.abs(points - o_points).max()),
            "pose_abs_vs_one_gpu": float(np.abs(poses - one.poses()).max()),
            "point_abs_vs_one_gpu": float(np.abs(points - one.points()).max()),
            "ranks_bit_identical": bool(flags.item() == 1.0),
            "device_ms": s.device_ms, "one_gpu_device_ms": s1.device_ms,
        }
        ok = (out["cost_rel_vs_oracle"] < 1e-5 and out["pose_abs_vs_oracle"] < 1e-6 and out["point_abs_vs_oracle"] < 1e-5
              and out["ranks_bit_identical"] and s.iterations == it)
        out["ok"] = ok
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
