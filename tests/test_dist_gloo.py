"""N > 1 host logic on CPU: two processes, gloo backend, 127.0.0.1 rendezvous.
Covers what surrounds the C-ABI call in paddle2_b200/dist.py: the row partition, the
unique-id hand-off from rank 0, the global row count, and -- with the oracle standing
in for each rank's local GPU reduce, as the checker only -- that per-shard partials
combined by a SUM allreduce equal the reduce of the whole tensor."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pyoracle as O
        from paddle2_b200 import dist as pdist
        uid = pdist.exchange_unique_id(lambda: bytes(range(128)))
        assert uid == bytes(range(128))
        total_rows, cols = 1003, 64
        start, count = pdist.partition_rows(total_rows, world, rank)
        assert pdist.global_count(count) == total_rows
        rng = np.random.default_rng(123)
        full = rng.random((total_rows, cols)).astype(np.float32)
        local = full[start:start + count]
        partial = torch.from_numpy(O.reduce("sum", local, [0]))   # stand-in for the GPU kernel
        dist.all_reduce(partial, op=dist.ReduceOp.SUM)
        want = O.reduce("sum", full, [0])
        np.testing.assert_allclose(partial.numpy(), want, rtol=1e-5)
        # max / min shard the same way with their own red-ops (gpu/all_reduce_kernel.cu:46-65)
        pm = torch.from_numpy(O.reduce("max", local, [0]))
        dist.all_reduce(pm, op=dist.ReduceOp.MAX)
        assert np.array_equal(pm.numpy(), O.reduce("max", full, [0]))
        # sharded mean = allreduced sum / GLOBAL count
        mean = partial.numpy() / np.float32(pdist.global_count(count))
        np.testing.assert_allclose(mean, O.reduce("mean", full, [0]), rtol=1e-5)
        # sharded argmax: per-rank (value, global index) pairs, all-gathered, lowest index on ties
        assert pdist.row_offset_of(count) == start
        ti = np.random.default_rng(7).integers(0, 4, size=(total_rows, cols)).astype(np.int64)
        lt = ti[start:start + count]
        li = O.argreduce("argmax", lt, 0) + pdist.row_offset_of(count)   # stand-in for the GPU kernel
        lv = O.reduce("max", lt, [0])
        pairs = [None] * world
        dist.all_gather_object(pairs, (lv, li))
        bv, bi = pairs[0]
        for v, i in pairs[1:]:
            take = (v > bv) | ((v == bv) & (i < bi))
            bv, bi = np.where(take, v, bv), np.where(take, i, bi)
        assert np.array_equal(bi, O.argreduce("argmax", ti, 0))
        # single-rank ShardedReducer never creates a communicator
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}
