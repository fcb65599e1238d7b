"""CPU: the N>1 host logic (block partition + the path's one collective, the count all-gather)
under world_size-2 gloo."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_partition_covers_batch():
    from isb.sharding import partition
    for total in (512, 64, 7, 1):
        for world in (1, 2, 4, 8):
            spans = [partition(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert partition(512, 8, 3) == (192, 256)


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from isb.sharding import gather_counts, partition
    all_counts = (torch.arange(total, dtype=torch.int32) * 7) % 101
    all_counts[3] = -1           # an overflow marker must survive the gather and count as 0 detections
    a, b = partition(total, world, rank)
    counts, offsets = gather_counts(all_counts[a:b])
    counts2, offsets2 = gather_counts(all_counts[a:b], pad_to=(total + world - 1) // world) if total % world == 0 else (counts, offsets)
    q.put((rank, counts.tolist(), offsets.tolist(), counts2.tolist(), offsets2.tolist()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [8, 7])
def test_count_gather_world2(total):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = ((torch.arange(total, dtype=torch.int64) * 7) % 101)
    expect[3] = -1
    valid = expect.clamp(min=0)
    off = (torch.cumsum(valid, 0) - valid).tolist()
    for rank, counts, offsets, counts2, offsets2 in res:
        assert counts == expect.tolist() and offsets == off
        assert counts2 == expect.tolist() and offsets2 == off


def test_single_process_fallback():
    from isb.sharding import gather_counts
    c, o = gather_counts(torch.tensor([3, 0, 5], dtype=torch.int32))
    assert c.tolist() == [3, 0, 5] and o.tolist() == [0, 3, 3]
