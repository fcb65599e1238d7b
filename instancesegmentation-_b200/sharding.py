"""Per-image sharding of the post-processing batch across the GPUs of one node.

Images are independent on this path (reference: batch-1 `squeeze()` at utils_bbox.py:177-180;
per-image loops at rpn_function.py:507, roi_head.py:413), so rank g owns a contiguous block of
the batch and no data-path collective exists.  The only exchange is one all-gather of the
per-image detection counts so that every rank can compute global output offsets; mask payloads
stay on the GPU that produced them.
"""
from __future__ import annotations

import ctypes as C

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


class PeerCountExchange:
    """The count exchange FUSED into the detect kernel over NVLink peer memory (include/isb.h `isb_counts_*`,
    csrc/counts_exchange.cu): after `attach`, every `decode_nms_batch` / `isb_yolact_detect` call of `n_local` images stores
    its counts straight into a ring in every rank's HBM from the tail of its NMS kernel -- no NCCL kernel competing with the
    persistent mask kernel for an SM, no cross-rank stream wait, nothing to post.  `result()` enqueues one small kernel that
    waits (bounded) for the latest epoch of all ranks and returns (counts, offsets) like `gather_counts`.

    Same interface as `CountExchange` (`post` / `flush` are no-ops here), so a caller can take whichever is available:
    `open_count_exchange` tries this one and falls back to the NCCL all-gather when CUDA IPC between the ranks' devices
    is not possible.  Create it AFTER the per-device handle has its final limits (growing the handle re-creates it and
    drops the ring), with the process group initialised; all ranks must issue the same sequence of detect calls."""

    fused = True

    def __init__(self, n_local, device, group=None, depth=4):
        from . import _lib
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.n_local, self.device = int(n_local), torch.device(device)
        self.handle = _lib.get_handle(self.device.index)
        mine = C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            self.handle.check(self.handle.lib.isb_counts_export(self.handle.ptr, self.n_local, self.world, int(depth), mine))
            everyone = [None] * self.world
            dist.all_gather_object(everyone, mine.raw, group=group)     # also orders every rank's export before any attach
            try:
                self.handle.check(self.handle.lib.isb_counts_attach(self.handle.ptr, self.rank, b"".join(everyone)))
                ok = 1
            except Exception as e:  # noqa: BLE001  (reported below, collectively)
                ok, self.error = 0, e
            flags = [None] * self.world
            dist.all_gather_object(flags, ok, group=group)
            if not all(flags):
                self.handle.lib.isb_counts_detach(self.handle.ptr)
                raise RuntimeError(f"CUDA IPC count exchange unavailable on ranks {[r for r, f in enumerate(flags) if not f]}"
                                   + (f": {self.error}" if not ok else ""))
        n = self.world * self.n_local
        self.counts = torch.zeros((n,), dtype=torch.int32, device=self.device)
        self.offsets = torch.zeros((n,), dtype=torch.int64, device=self.device)
        self.status = torch.zeros((1,), dtype=torch.int32, device=self.device)

    def post(self, local_counts=None):
        return int(self.handle.lib.isb_counts_epoch(self.handle.ptr))

    def flush(self):
        pass

    def result(self, epoch=-1, timeout_ms=2000):
        """(counts int64 [world * n_local], offsets int64) of `epoch` (-1: this rank's latest detect call), on the current
        stream; `self.status` (device int32) is 0 unless a peer never published (1) or the slot was overwritten (2)."""
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            self.handle.check(self.handle.lib.isb_counts_wait(self.handle.ptr, int(epoch), C.c_void_p(self.counts.data_ptr()),
                                                              C.c_void_p(self.offsets.data_ptr()), C.c_void_p(self.status.data_ptr()),
                                                              int(timeout_ms), stream))
        return self.counts.to(torch.int64), self.offsets

    def close(self):
        """Collective: every rank stops publishing, then the rings are unmapped and freed."""
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        self.handle.lib.isb_counts_detach(self.handle.ptr)


def open_count_exchange(n_local, device, group=None, prefer_fused=True):
    """The count exchange for a sharded job with equal shards: the fused NVLink peer-store version when the ranks' devices
    can map each other's memory, else the NCCL all-gather (`CountExchange`)."""
    if prefer_fused and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1 \
            and torch.device(device).type == "cuda":
        try:
            return PeerCountExchange(n_local, device, group)
        except Exception:  # noqa: BLE001  (every rank raises together: the decision is collective)
            pass
    return CountExchange(n_local, device, group)
