"""Where the end-to-end time of a SHARDED large problem goes (BASELINE configs 3 and 4), per rank: set_problem (staging),
upload (structure build), LM, read-back — with the caller's arrays in page-locked memory.  Run under torchrun."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import viso_b200 as vb

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
for cid in (3, 4):
    p = vb.synth_config(cid)
    pin = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in p.items() if isinstance(v, np.ndarray)}
    a = {k: t.numpy() for k, t in pin.items()}
    ba = vb.BundleAdjuster(len(a["points_init"]), len(a["poses_init"]), len(a["edge_pose"]), device=local, huber_delta=p["huber_delta"])
    ident = [vb.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ident, src=0)
    ba.shard_init(rank, world, ident[0])
    for rep in range(3):
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ba.set_problem(a["poses_init"], a["points_init"], a["edge_pose"], a["edge_point"], a["edge_uv"])
        t1 = time.perf_counter()
        ba.upload()
        t2 = time.perf_counter()
        s = ba.optimize_resident(10)
        t3 = time.perf_counter()
        ba.download()
        t4 = time.perf_counter()
        print("rank %d config %d rep %d: set_problem %.1f ms, upload %.1f ms, LM %.1f ms (device %.1f), download %.1f ms, total %.1f ms" % (
            rank, cid, rep, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), s.device_ms, 1e3 * (t4 - t3), 1e3 * (t4 - t0)), flush=True)
    ba.close()
dist.destroy_process_group()
