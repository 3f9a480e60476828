#!/usr/bin/env python
"""host-copy ceiling of the box: every rank copies pinned host memory to its GPU and back at the same time, in the
proportions of one C2 e2e call (1 GiB in, 0.45 GiB out) on two streams; prints the aggregate "text GB/s" those copies
alone would allow.  Run under torchrun with as many ranks as GPUs (or alone for one GPU)."""
import os, time
import torch
import torch.distributed as dist
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n_in, n_out = 1 << 30, int(0.45 * (1 << 30))
h_in = torch.empty(n_in, dtype=torch.uint8).pin_memory(); h_out = torch.empty(n_out, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n_in, dtype=torch.uint8, device="cuda"); d_out = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def once():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)
for _ in range(2):
    once()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
K = 5
for _ in range(K):
    once()
torch.cuda.synchronize()
dt = torch.tensor([time.perf_counter() - t0], device="cuda")
if world > 1:
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
if rank == 0:
    sec = float(dt.item()) / K
    print("copy ceiling: %d rank(s), %.1f ms per (1 GiB in + 0.45 GiB out) per rank -> %.1f GB/s of text aggregate (%.1f per GPU)" % (
        world, sec * 1e3, world * n_in / sec / 1e9, n_in / sec / 1e9))
