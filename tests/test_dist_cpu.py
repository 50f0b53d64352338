"""CPU: the multi-rank host logic (sharding + record all-gather) with the gloo backend, world_size 2 and 3."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fairify_b200.dist import allgather_records, pack_records, shard_range, unpack_records


def test_shard_range_is_a_partition_of_the_index_space():
    for n in (0, 1, 7, 16000, 3290112, 10 ** 7 + 3):
        for world in (1, 2, 3, 4, 8):
            nxt = 0
            for r in range(world):
                first, count = shard_range(n, r, world)
                assert first == nxt and count >= 0
                nxt += count
            assert nxt == n
            sizes = [shard_range(n, r, world)[1] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def _fake_records(first, count, NW):
    idx = torch.arange(first, first + count, dtype=torch.int64)
    verdict = (idx % 5).to(torch.uint8)
    dead = torch.stack([(idx * (k + 3) % 65521).to(torch.int32) for k in range(NW)], dim=1)
    cand = torch.stack([(idx * (k + 11) % 32749).to(torch.int32) for k in range(NW)], dim=1)
    ce = ((idx * 7) % 1001 - 1).to(torch.int32)
    return verdict, dead, cand, ce


def _worker(rank, world, port, n_total, NW, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        first, count = shard_range(n_total, rank, world)
        v, d, c, e = _fake_records(first, count, NW)
        full = allgather_records(pack_records(v, d, c, e), n_total)
        gv, gd, gc, ge = unpack_records(full, NW)
        ev, ed, ec, ee = _fake_records(0, n_total, NW)
        ok = bool(torch.equal(gv, ev) and torch.equal(gd, ed) and torch.equal(gc, ec) and torch.equal(ge, ee))
        ret[rank] = ok
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,n_total,NW", [(2, 1001, 1), (2, 64, 10), (3, 100, 2)])
def test_allgather_records_gloo(world, n_total, NW):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n_total, NW, ret), nprocs=world, join=True)
    assert all(ret.get(r) for r in range(world)), dict(ret)
