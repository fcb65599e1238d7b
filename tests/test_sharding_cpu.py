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


def _xchg_worker(rank, world, port, n_local, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from isb.sharding import CountExchange
    x = CountExchange(n_local, torch.device("cpu"), depth=2)
    got = []
    for step in range(5):   # more posts than slots: the oldest in-flight gather is waited for and its slot reused
        local = (torch.arange(n_local, dtype=torch.int32) + 10 * rank + 100 * step)
        if step == 4 and rank == 1:
            local[0] = -1
        x.post(local)
        if step in (1, 4):
            c, o = x.result()
            got.append((c.tolist(), o.tolist()))
    x.flush()
    q.put((rank, got))
    dist.barrier()
    dist.destroy_process_group()


def test_count_exchange_world2():
    world, port, n_local = 2, _free_port(), 4
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_xchg_worker, args=(r, world, port, n_local, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for step_i, step in enumerate((1, 4)):
        expect = torch.cat([torch.arange(n_local, dtype=torch.int64) + 10 * r + 100 * step for r in range(world)])
        if step == 4:
            expect[n_local] = -1
        valid = expect.clamp(min=0)
        off = (torch.cumsum(valid, 0) - valid).tolist()
        for rank, got in res:
            assert got[step_i] == (expect.tolist(), off)


def test_count_exchange_single_process():
    from isb.sharding import CountExchange
    x = CountExchange(3, torch.device("cpu"))
    x.post(torch.tensor([3, 0, 5], dtype=torch.int32))
    c, o = x.result()
    assert c.tolist() == [3, 0, 5] and o.tolist() == [0, 3, 3]


def test_open_count_exchange_falls_back_without_cuda():
    """open_count_exchange picks the fused NVLink version only for CUDA devices of an initialised multi-rank group; anywhere
    else it is the all-gather class (single process here: world 1, CPU tensors)."""
    from isb.sharding import CountExchange, open_count_exchange
    x = open_count_exchange(4, torch.device("cpu"))
    assert isinstance(x, CountExchange) and not getattr(x, "fused", False)
    x.post(torch.tensor([3, 0, -1, 5], dtype=torch.int32))
    counts, offsets = x.result()
    assert counts.tolist() == [3, 0, -1, 5] and offsets.tolist() == [0, 3, 3, 3]
