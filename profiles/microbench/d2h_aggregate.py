"""Aggregate device->pinned-host copy bandwidth with one process per GPU, all copying at once: the
ceiling of bench.py's end-to-end figure, whose ranks each send 5.9 GB of checkpoints per 0.39 s step.
    python -m torch.distributed.run --nproc-per-node N profiles/microbench/d2h_aggregate.py"""
import os
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 27                                             # 1 GiB of fp64
src = torch.ones(n, dtype=torch.float64, device="cuda")
dst = torch.empty(n, dtype=torch.float64).pin_memory()
for alone in ((False, True) if world > 1 else (True,)):
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 6
    if not alone or rank == 0:
        e0.record()
        for _ in range(reps):
            dst.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        gbs = reps * n * 8 / (e0.elapsed_time(e1) * 1e-3) / 1e9
    else:
        gbs = 0.0
    t = torch.tensor([gbs], device="cuda")
    if world > 1:
        dist.barrier()
        dist.all_reduce(t)
    if rank == 0:
        print("%d GPU(s) copying %s: %.1f GB/s aggregate (rank 0: %.1f GB/s)"
              % (1 if alone else world, "alone" if alone else "at once", float(t), gbs), flush=True)
if world > 1:
    dist.destroy_process_group()
