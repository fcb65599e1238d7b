"""Per-image sharding of the post-processing batch across the GPUs of one node.

Images are independent on this path (reference: batch-1 `squeeze()` at utils_bbox.py:177-180;
per-image loops at rpn_function.py:507, roi_head.py:413), so rank g owns a contiguous block of
the batch and no data-path collective exists.  The only exchange is one all-gather of the
per-image detection counts so that every rank can compute global output offsets; mask payloads
stay on the GPU that produced them.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def partition(total, world_size, rank):
    """Contiguous block partition: rank r owns images [start, stop) of `total`."""
    base, rem = divmod(int(total), int(world_size))
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def gather_counts(local_counts, group=None, pad_to=None):
    """All-gather per-image detection counts (int32 [B_local]) -> (counts [B_total], offsets
    [B_total]) identical on every rank; offsets = exclusive scan, i.e. where each image's
    detections start in a concatenated result.  Works for unequal shard sizes via padding."""
    if not (dist.is_available() and dist.is_initialized()):
        counts = local_counts.to(torch.int64)
        return counts, torch.cumsum(counts, 0) - counts
    world = dist.get_world_size(group)
    n_local = int(local_counts.shape[0])
    if pad_to is None:
        sizes = torch.tensor([n_local], dtype=torch.int64, device=local_counts.device)
        all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
        dist.all_gather(all_sizes, sizes, group=group)
        all_sizes = [int(s.item()) for s in all_sizes]
        pad_to = max(all_sizes)
    else:
        all_sizes = None
    buf = torch.full((pad_to,), -2, dtype=torch.int32, device=local_counts.device)
    buf[:n_local] = local_counts.to(torch.int32)
    out = torch.empty((world * pad_to,), dtype=torch.int32, device=local_counts.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    out = out.view(world, pad_to)
    if all_sizes is None:
        counts = out.reshape(-1)
        counts = counts[counts != -2]
    else:
        counts = torch.cat([out[r, : all_sizes[r]] for r in range(world)])
    counts = counts.to(torch.int64)
    valid = counts.clamp(min=0)
    return counts, torch.cumsum(valid, 0) - valid


class CountExchange:
    """The same collective for the steady state of a sharded job with equal shards, taken off the critical path.

    `post(local_counts)` snapshots this step's counts (one tiny device copy on the caller's stream) and starts the
    all-gather asynchronously: NCCL runs it on its own stream behind that copy, so it overlaps the next step's
    score-filter kernel instead of adding its latency to every step, and nothing synchronises with the host.
    `result()` makes the caller's stream wait for the most recent gather and returns (counts, offsets) like
    `gather_counts`.  `depth` snapshots are kept in flight; posting the (depth+1)-th waits for the oldest.
    """

    def __init__(self, n_local, device, group=None, depth=2):
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.n_local = int(n_local)
        self.stage = [torch.zeros((self.n_local,), dtype=torch.int32, device=device) for _ in range(depth)]
        self.gathered = [torch.zeros((self.world * self.n_local,), dtype=torch.int32, device=device) for _ in range(depth)]
        self.work = [None] * depth
        self.posted = 0

    def post(self, local_counts):
        slot = self.posted % len(self.stage)
        if self.work[slot] is not None:
            self.work[slot].wait()  # stream-level wait on NCCL (host-blocking only on gloo)
            self.work[slot] = None
        self.stage[slot].copy_(local_counts)
        if self.world > 1:
            self.work[slot] = dist.all_gather_into_tensor(self.gathered[slot], self.stage[slot], group=self.group, async_op=True)
        else:
            self.gathered[slot].copy_(self.stage[slot])
        self.posted += 1
        return slot

    def flush(self):
        for i, w in enumerate(self.work):
            if w is not None:
                w.wait()
                self.work[i] = None

    def result(self):
        if self.posted == 0:
            raise RuntimeError("CountExchange.result() before any post()")
        slot = (self.posted - 1) % len(self.stage)
        if self.work[slot] is not None:
            self.work[slot].wait()
            self.work[slot] = None
        counts = self.gathered[slot].to(torch.int64)
        valid = counts.clamp(min=0)
        return counts, torch.cumsum(valid, 0) - valid
