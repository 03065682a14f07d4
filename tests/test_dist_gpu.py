"""Two-rank NCCL run of the sharded reduce (skipped unless >= 2 GPUs are visible):
one process per GPU, local reduce + allreduce in one C-ABI call."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


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
    sys.path.insert(0, os.path.join(root, "tests"))
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pyoracle as O
        from paddle2_b200 import _lib as L
        from paddle2_b200.dist import ShardedReducer, partition_rows
        total_rows, cols = 40003, 1024
        full = np.random.default_rng(123).random((total_rows, cols)).astype(np.float32)
        start, count = partition_rows(total_rows, world, rank)
        x = torch.from_numpy(full[start:start + count]).cuda()
        red = ShardedReducer()
        import gpu_util as G
        got = red.sum(x).cpu().numpy()
        # the reference adds the 40003 rows one after another in fp32: bound widened by its own error
        G.assert_reduce_close(got, O.reduce("sum", full, [0]), full, [0], "sum", G.FP32_RTOL)
        got = red.reduce(L.MAX, x).cpu().numpy()
        assert np.array_equal(got, O.reduce("max", full, [0]))
        got = red.reduce(L.MEAN, x).cpu().numpy()
        G.assert_reduce_close(got, O.reduce("mean", full, [0]), full, [0], "mean", G.FP32_RTOL)
        torch.cuda.synchronize()
        # bf16 sum: the ranks' partials travel in fp32 and the result is rounded once
        import paddle2_b200 as pb
        xb16 = x.bfloat16()
        part = pb.sum(xb16, axis=0, dtype="float32").cpu().numpy()
        parts = [None] * world
        dist.all_gather_object(parts, part)
        want16 = torch.from_numpy(np.sum(np.stack(parts), axis=0, dtype=np.float32)).bfloat16()
        assert torch.equal(red.sum(xb16).cpu(), want16)
        got16 = red.reduce(L.MEAN, xb16).float().cpu().numpy()
        np.testing.assert_allclose(got16, np.stack(parts).sum(0) / total_rows, rtol=G.BF16_RTOL)
        # without a collective library: exchange kernel over peer memory behind the local reduce
        from paddle2_b200.dist import P2PReducer
        p2p = P2PReducer(max_out_numel=cols, max_tiles=64)
        assert torch.equal(p2p.sum(xb16).cpu(), want16)
        nccl_sum = red.sum(x)
        for _ in range(3):                       # epochs 1..3: slots alternate by parity
            fused = p2p.sum(x)
            torch.cuda.synchronize()
            np.testing.assert_allclose(fused.cpu().numpy(), nccl_sum.cpu().numpy(), rtol=1e-6)
        G.assert_reduce_close(fused.cpu().numpy(), O.reduce("sum", full, [0]), full, [0], "sum", G.FP32_RTOL)
        assert np.array_equal(p2p.reduce(L.MAX, x).cpu().numpy(), O.reduce("max", full, [0]))
        fm = p2p.reduce(L.MEAN, x, global_rows=total_rows).cpu().numpy()
        G.assert_reduce_close(fm, O.reduce("mean", full, [0]), full, [0], "mean", G.FP32_RTOL)
        # every rank must hold the same bits: compare through a gloo all_gather
        gathered = [None] * world
        dist.all_gather_object(gathered, fused.cpu().numpy().tobytes())
        assert all(b == gathered[0] for b in gathered)
        p2p.status()
        p2p.close()
        # argmax / argmin across the shards: ties span ranks, the lowest GLOBAL row must win
        ti = np.random.default_rng(7).integers(0, 5, size=(total_rows, cols)).astype(np.int64)
        xs = torch.from_numpy(ti[start:start + count]).cuda()
        gi, gv = red.argmax(xs, want_value=True)
        assert np.array_equal(gi.cpu().numpy(), O.argreduce("argmax", ti, 0))
        assert np.array_equal(gv.cpu().numpy(), ti.max(0))
        gi = red.argmin(x, dtype="int32")
        assert gi.dtype == torch.int32
        assert np.array_equal(gi.cpu().numpy(), O.argreduce("argmin", full, 0, dtype=np.int32))
        xb = torch.from_numpy(full[start:start + count]).cuda().bfloat16()     # bf16: natural ties
        fb = torch.from_numpy(full).bfloat16().float().numpy()
        assert np.array_equal(red.argmax(xb).cpu().numpy(), O.argreduce("argmax", fb, 0))
        red.close()
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_two_rank_nccl_sharded_sum():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}
