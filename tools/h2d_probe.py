"""Host -> device bandwidth of page-locked memory with ONE rank copying and with ALL ranks copying at once (torchrun):
what bounds the end-to-end C5 step at N = 8 (every rank uploads 296 MB of edges per step)."""
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
n = 296 * 1024 * 1024
src = torch.empty(n, dtype=torch.uint8).pin_memory()
src.fill_(1)
dst = torch.empty(n, dtype=torch.uint8, device="cuda")
back = torch.empty(48 * 1024 * 1024, dtype=torch.uint8).pin_memory()


def run(active, reps=20):
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if active:
        for _ in range(reps):
            dst.copy_(src, non_blocking=True)
            back.copy_(dst[: back.numel()], non_blocking=True)
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    dist.barrier()
    return (n * reps / dt / 1e9) if active else 0.0


run(True, 3)
alone = run(rank == 0)
together = run(True)
t = torch.tensor([alone, together], dtype=torch.float64, device="cuda")
lst = [torch.zeros_like(t) for _ in range(world)]
dist.all_gather(lst, t)
if rank == 0:
    tg = [float(x[1]) for x in lst]
    print("H2D of 296 MB pinned (+ 48 MB D2H) per step: rank 0 alone %.1f GB/s; all %d ranks at once: per rank min %.1f / mean %.1f GB/s, aggregate %.1f GB/s"
          % (float(lst[0][0]), world, min(tg), sum(tg) / world, sum(tg)), flush=True)
dist.destroy_process_group()
