"""Host<->device copy bandwidth with all ranks copying at once (developer tool):
torchrun --nproc-per-node N tools/pcie_probe.py"""
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(rank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
n = 128 * 1024 * 1024 // 8
h = torch.empty(n, dtype=torch.float64).pin_memory()
h2 = torch.empty(n, dtype=torch.float64).pin_memory()
d = torch.empty(n, dtype=torch.float64, device="cuda")
d2 = torch.ones(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ("h2d", "d2h", "both"):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s1):
                d.copy_(h, non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s2):
                h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gb = 10 * n * 8 / 1e9 * (2 if mode == "both" else 1)
    print(f"rank {rank}/{world} {mode}: {gb / dt:.1f} GB/s", flush=True)
if world > 1:
    dist.destroy_process_group()
