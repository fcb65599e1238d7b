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
