"""N>1 host logic on CPU (gloo, world_size 2): the point partition of viso_ba_shard_init
(viso_ba_shard_point_ranges) and the one exchange step of the sharded path — every rank
linearises and Schur-reduces ITS points only, the partial reduced camera systems and costs are
summed with one fp64 allreduce, and only rank 0 contributes the damping of the pose diagonal
(DevProblem::lambda_scale).  The oracle plays the per-rank kernels; the sum must equal the
unsharded reduced system.  The NCCL version of the same exchange is tests/test_gpu_sharded.py (`-m gpu`, two
ranks under torch.distributed.run, skipped on a one-GPU box) and the `sharded` section of bench.py at N > 1."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    import oracle_lib as orc
    import viso_b200 as vb
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        p = vb.synth_config(1, n_poses=7, n_points=300, target_edges=1500, seed=0x5EED0077)
        lam = 0.37
        c = orc.cfg(K=p["K"], huber_delta=2.0)
        bounds = vb.shard_point_ranges(len(p["points_init"]), p["edge_point"], world)
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        mine = (p["edge_point"] >= lo) & (p["edge_point"] < hi)
        # this rank's share: its edges only (the other ranks' points then have no edges and add nothing)
        S, bs, _ = orc.reduced_system(c, p["poses_init"], p["points_init"], p["edge_pose"][mine], p["edge_point"][mine],
                                      p["edge_uv"][mine], lam)
        if rank != 0:
            S -= lam * np.eye(len(S))  # lambda_scale = 0 off rank 0
        cost = orc.cost(c, p["poses_init"], p["points_init"], p["edge_pose"][mine], p["edge_point"][mine], p["edge_uv"][mine])
        buf = torch.from_numpy(np.concatenate([S.ravel(), bs, [cost]]))
        dist.all_reduce(buf, op=dist.ReduceOp.SUM)  # the path's only data exchange per linearisation
        n = len(bs)
        S_all, bs_all, cost_all = buf[: n * n].numpy().reshape(n, n), buf[n * n: n * n + n].numpy(), float(buf[-1])
        S_ref, bs_ref, _ = orc.reduced_system(c, p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"], lam)
        cost_ref = orc.cost(c, p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
        scale = np.abs(S_ref).max()
        assert np.abs(S_all - S_ref).max() <= 1e-12 * scale
        assert np.abs(bs_all - bs_ref).max() <= 1e-12 * np.abs(bs_ref).max()
        assert abs(cost_all - cost_ref) <= 1e-12 * cost_ref
        # the partition: contiguous, covers every point once, balanced on edges
        assert bounds[0] == 0 and bounds[-1] == len(p["points_init"]) and np.all(np.diff(bounds) >= 0)
        share = mine.sum() / len(p["edge_point"])
        assert abs(share - 1.0 / world) < 0.05
        open(os.path.join(out_dir, "ok%d" % rank), "w").write("%d %d" % (lo, hi))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_sharded_reduced_system_sums_to_the_unsharded_one(vb, oracle, tmp_path, world):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(os.path.join(str(tmp_path), "ok%d" % r)) for r in range(world))


def test_shard_ranges_edge_cases(vb):
    ep = np.array([0, 0, 0, 0, 3, 3], np.int32)
    b = vb.shard_point_ranges(5, ep, 3)
    assert b[0] == 0 and b[-1] == 5 and np.all(np.diff(b) >= 0)
    assert list(vb.shard_point_ranges(4, np.zeros(0, np.int32), 2)) [0] == 0
    with pytest.raises(vb.VisoBaError):
        vb.shard_point_ranges(2, np.array([5], np.int32), 2)
