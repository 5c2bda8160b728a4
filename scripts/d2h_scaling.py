#!/usr/bin/env python
"""What the box's host side gives N GPUs at once: plain cudaMemcpyAsync device->host of a 178 MB buffer (the size of one
4096-chunk terrain batch's vertex streams) into pinned host memory, every rank at the same time, no mesher involved.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/d2h_scaling.py
Prints one JSON line: per-rank and whole-box GB/s -- the ceiling of the bench's e2e number at N GPUs."""
import json
import os
import time

import torch

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nbytes, reps = 178_608_216, 30
src = torch.empty(nbytes, dtype=torch.uint8, device="cuda").fill_(3)
dst = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
for _ in range(3):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.time()
for _ in range(reps):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
dt = time.time() - t0
gbs = nbytes * reps / dt / 1e9
if world > 1:
    t = torch.tensor([gbs, dt], dtype=torch.float64, device="cuda")
    allr = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allr, t)
    per = [float(x[0]) for x in allr]
    slow = max(float(x[1]) for x in allr)
else:
    per, slow = [gbs], dt
if rank == 0:
    print(json.dumps({"n_gpus": world, "bytes_per_copy": nbytes, "per_rank_gbs": [round(v, 1) for v in per],
                      "whole_box_gbs": round(world * nbytes * reps / slow / 1e9, 1), "cpus": os.cpu_count()}))
if world > 1:
    dist.destroy_process_group()
