"""N > 1 protocol on CPU (gloo, world_size 2 and 3): train rows partitioned by
spectavi_b200.sharding.shard_range, each rank's top-2 packed as the device key
format (dist << 32 | global idx, all-ones = none), all-gathered, merged
lexicographically -- must equal the single-rank oracle bit for bit.  The per-rank
matcher here is the CPU oracle (test infrastructure); on the GPU the same
protocol runs with K1 / K5 (tests/test_multigpu.py)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def pack_keys(idx, dist, offset):
    keys = (dist.astype(np.uint64) << np.uint64(32)) | (idx + np.uint64(offset))
    keys[idx == NONE] = NONE
    return np.ascontiguousarray(keys.T)            # [2][yrows]


def merge_keys(gathered):
    """[world][2][yrows] -> [2][yrows]: two smallest distinct keys per query"""
    allk = np.concatenate(list(gathered), axis=0)  # [2 * world][yrows]
    allk = np.sort(allk, axis=0)
    return allk[:2]


def unpack_keys(keys):
    idx = (keys & np.uint64(0xFFFFFFFF)).astype(np.uint64)
    dist = (keys >> np.uint64(32)).astype(np.int64)
    none = keys == NONE
    idx[none] = NONE
    dist[none] = 2**31 - 1
    return idx.T, dist.T.astype(np.int32)


def _worker(rank, world, port, xrows, yrows, dim, out):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import pyoracle as po
    from spectavi_b200.sharding import shard_range
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    rng = np.random.RandomState(1234)              # same data on every rank
    x = rng.randint(0, 4, (xrows, dim)).astype("uint8")     # small alphabet: ties across shards
    y = rng.randint(0, 4, (yrows, dim)).astype("uint8")
    x0, xn = shard_range(xrows, world, rank)
    idx, d = po.l1k2(np.ascontiguousarray(x[x0:x0 + xn]), y, 1)
    keys = torch.from_numpy(pack_keys(idx, d, x0).view(np.int64))
    parts = [torch.empty_like(keys) for _ in range(world)]
    dist.all_gather(parts, keys)
    merged = merge_keys(torch.stack(parts).numpy().view(np.uint64))
    gi, gd = unpack_keys(merged)
    fi, fd = po.l1k2(x, y, 1)
    ok = bool((gi == fi).all() and (gd == fd).all())
    res = torch.tensor([int(ok)])
    dist.all_reduce(res, op=dist.ReduceOp.MIN)
    if rank == 0:
        out.put(int(res.item()))
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,xrows", [(2, 301), (3, 100), (2, 1), (3, 2)])
def test_sharded_protocol_gloo(world, xrows):
    import torch.multiprocessing as mp
    from oracle import pyoracle as po
    po.lib()                                       # build the checker once, before forking
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, xrows, 37, 32, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) == 1


def test_shard_range_covers_everything():
    from spectavi_b200.sharding import shard_range
    for xrows in (0, 1, 2, 7, 8, 1000, 1_000_000):
        for world in (1, 2, 3, 4, 8):
            pos = 0
            for r in range(world):
                x0, xn = shard_range(xrows, world, r)
                assert x0 == min(pos, xrows) and xn >= 0
                pos = x0 + xn
            assert pos == xrows
