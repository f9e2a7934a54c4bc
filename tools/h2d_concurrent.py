#!/usr/bin/env python
"""How fast can N ranks copy pinned host memory to their GPUs AT THE SAME TIME?  (bench.py's e2e feeds 208 MB per rank per
step; SCALE_r01 showed its per-rank time growing from 3.85 ms at N <= 2 to 9.0 ms at N = 8.)  Each rank copies the same
208,142,336-byte pinned buffer 20 times after a barrier; rank 0 prints per-rank and aggregate GB/s.  No kernels run.
usage: torchrun --nproc-per-node N tools/h2d_concurrent.py"""
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
nbytes = 65536 * 3176
host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
host.fill_(rank + 1)
devbuf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
for _ in range(3):
    devbuf.copy_(host, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    devbuf.copy_(host, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
gbs = torch.tensor([20 * nbytes / dt / 1e9], dtype=torch.float64, device=dev)
allg = [torch.zeros_like(gbs) for _ in range(world)]
if world > 1:
    dist.all_gather(allg, gbs)
else:
    allg = [gbs]
if rank == 0:
    vals = [float(g.item()) for g in allg]
    print("concurrent pinned H2D, %d ranks, %d bytes x 20: per rank GB/s %s | aggregate %.1f GB/s | ms per 208 MB copy %.2f (slowest rank)"
          % (world, nbytes, " ".join("%.1f" % v for v in vals), sum(vals), nbytes / min(vals) / 1e6))
if world > 1:
    dist.destroy_process_group()
