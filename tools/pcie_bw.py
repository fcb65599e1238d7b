"""PCIe / host-memory ceiling for the e2e path: pinned D2H, H2D and both at once (two streams) -- on ONE GPU, or on all ranks
of a torchrun launch CONCURRENTLY (every rank copies at the same time between barriers), which shows what the box's host side
sustains in aggregate:   python -m torch.distributed.run --nproc-per-node N tools/pcie_bw.py"""
import json
import os
import time

import torch

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
n = 1 << 30
d = torch.empty(n, dtype=torch.uint8, device="cuda")
hdst = torch.empty(n, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
hsrc = torch.empty(n, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(do_d2h, do_h2d, reps=5):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if do_d2h:
            with torch.cuda.stream(s1):
                hdst.copy_(d, non_blocking=True)
        if do_h2d:
            with torch.cuda.stream(s2):
                d2.copy_(hsrc, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    return (time.perf_counter() - t0) / reps


run(True, True, 2)
r = {"ranks_copying_concurrently": world, "cpus": os.cpu_count(),
     "d2h_GBs_per_rank": n / run(True, False) / 1e9, "h2d_GBs_per_rank": n / run(False, True) / 1e9}
r["both_each_GBs_per_rank"] = n / run(True, True) / 1e9
for k in ("d2h", "h2d", "both_each"):
    r[f"{k}_GBs_aggregate"] = r[f"{k}_GBs_per_rank"] * world   # barrier-bracketed: the slowest rank sets the time
if rank == 0:
    print(json.dumps(r))
if world > 1:
    dist.destroy_process_group()
