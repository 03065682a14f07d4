"""b200phi_sharded_reduce_p2p exercised on ONE GPU: two "ranks" whose exchange buffers both
live on this device (no IPC mapping needed -- base[r] is simply the other buffer), their calls
issued on two streams so that both exchange kernels are resident together.  The peer stores,
flags, waits and the rank-ordered combine are exactly the code the 2..8-GPU path runs; what a
one-GPU box cannot show is NVLink itself (tests/test_dist_gpu.py does, on >= 2 GPUs).

Also the failure mode: a peer that never publishes must give PEER_TIMEOUT and an untouched
output, never a combine of stale slots."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import gpu_util as G
from paddle2_b200 import _lib as L

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _Pair:
    """Two ranks in one process on one device."""

    def __init__(self, max_out, max_tiles=64):
        self.lib = L.lib()
        n = self.lib.b200phi_p2p_buffer_bytes(2, max_out, max_tiles)
        assert n > 0
        self.bufs = []
        for _ in range(2):
            p = C.c_void_p()
            h = C.create_string_buffer(64)
            L.check(self.lib.b200phi_p2p_alloc(n, C.byref(p), h))
            self.bufs.append(p)
        self.groups = []
        for r in range(2):
            g = L.P2PGroup()
            g.nranks, g.rank = 2, r
            g.max_out_numel, g.max_tiles = max_out, max_tiles
            g.base[0], g.base[1] = self.bufs[0].value, self.bufs[1].value
            self.groups.append(g)
        self.streams = [torch.cuda.Stream(), torch.cuda.Stream()]
        self.epoch = 0

    def close(self):
        torch.cuda.synchronize()
        for p in self.bufs:
            self.lib.b200phi_p2p_free(p)

    def desc(self, x, out, op):
        d = L.ReduceDesc()
        d.op = op
        d.in_dtype = G_DT[x.dtype]
        d.out_dtype = G_DT[out.dtype]
        d.rank = x.dim()
        for i, s in enumerate(x.shape):
            d.dims[i] = s
        d.naxes = 1
        d.axes[0] = 0
        d.x, d.out = x.data_ptr(), out.data_ptr()
        ws_bytes = self.lib.b200phi_reduce_workspace(C.byref(d))
        ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device="cuda")
        d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
        return d, ws

    def run(self, shards, outs, op, global_rows=0, only_rank=None):
        self.epoch += 1
        keep = []
        torch.cuda.synchronize()
        rcs = []
        for r in range(2):
            if only_rank is not None and r != only_rank:
                continue
            d, ws = self.desc(shards[r], outs[r], op)
            keep.append(ws)
            rcs.append(self.lib.b200phi_sharded_reduce_p2p(C.byref(d), global_rows, C.byref(self.groups[r]),
                                                           self.epoch, C.c_void_p(self.streams[r].cuda_stream)))
        torch.cuda.synchronize()
        return rcs


G_DT = {torch.float32: L.FLOAT32, torch.bfloat16: L.BFLOAT16, torch.float16: L.FLOAT16,
        torch.int64: L.INT64, torch.int32: L.INT32, torch.float64: L.FLOAT64, torch.bool: L.BOOL}


def test_two_ranks_on_one_device_sum_max_mean(oracle):
    rows, cols = 30011, 1536
    full = np.random.default_rng(123).random((rows, cols)).astype(np.float32)
    cut = 14007
    shards = [torch.from_numpy(full[:cut]).cuda(), torch.from_numpy(full[cut:]).cuda()]
    pair = _Pair(max_out=cols)
    try:
        for _ in range(3):                                   # epochs 1..3: slots alternate by parity
            outs = [torch.full((cols,), float("nan"), device="cuda") for _ in range(2)]
            assert pair.run(shards, outs, L.SUM) == [0, 0], L.lib().b200phi_last_error()
            assert torch.equal(outs[0].view(torch.int32), outs[1].view(torch.int32))   # identical bits
            G.assert_reduce_close(G.to_np(outs[0]), oracle.reduce("sum", full, [0]), full, [0], "sum", G.FP32_RTOL)
        outs = [torch.empty(cols, device="cuda") for _ in range(2)]
        assert pair.run(shards, outs, L.MAX) == [0, 0]
        assert np.array_equal(G.to_np(outs[1]), oracle.reduce("max", full, [0]))
        assert pair.run(shards, outs, L.MEAN, global_rows=rows) == [0, 0]
        G.assert_reduce_close(G.to_np(outs[0]), oracle.reduce("mean", full, [0]), full, [0], "mean", G.FP32_RTOL)
        assert torch.equal(outs[0], outs[1])
        for p in pair.bufs:
            assert L.lib().b200phi_p2p_status(p) == 0
    finally:
        pair.close()


def test_sixteen_bit_sum_is_rounded_once():
    """bf16 / fp16 sums travel as fp32 partials: result == round(p0 + p1) of the fp32 partials."""
    import paddle2_b200 as pb
    g = torch.Generator(device="cuda").manual_seed(5)
    for dt in (torch.bfloat16, torch.float16):
        shards = [torch.rand(9001, 2048, device="cuda", generator=g).to(dt),
                  torch.rand(7003, 2048, device="cuda", generator=g).to(dt)]
        pair = _Pair(max_out=2048)
        try:
            outs = [torch.empty(2048, device="cuda", dtype=dt) for _ in range(2)]
            assert pair.run(shards, outs, L.SUM) == [0, 0], L.lib().b200phi_last_error()
            p0 = pb.sum(shards[0], axis=0, dtype="float32")
            p1 = pb.sum(shards[1], axis=0, dtype="float32")
            assert torch.equal(outs[0], (p0 + p1).to(dt))
            assert torch.equal(outs[0], outs[1])
            # a 16-bit exchange would have rounded each partial first: show the difference exists
            twice = (p0.to(dt).float() + p1.to(dt).float()).to(dt)
            assert not torch.equal(twice, outs[0])
        finally:
            pair.close()


def test_integer_sum_any_and_other_shapes(oracle):
    rng = np.random.default_rng(11)
    full = rng.integers(-1000, 1000, size=(4000, 7, 33)).astype(np.int64)
    shards = [torch.from_numpy(full[:1000]).cuda(), torch.from_numpy(full[1000:]).cuda()]
    pair = _Pair(max_out=7 * 33)
    try:
        outs = [torch.empty(7, 33, device="cuda", dtype=torch.int64) for _ in range(2)]
        assert pair.run(shards, outs, L.SUM) == [0, 0], L.lib().b200phi_last_error()
        assert np.array_equal(G.to_np(outs[0]), full.sum(0)) and torch.equal(outs[0], outs[1])
        fb = (rng.random((4000, 7, 33)) > 0.9995)
        sb = [torch.from_numpy(fb[:1000]).cuda(), torch.from_numpy(fb[1000:]).cuda()]
        ob = [torch.empty(7, 33, device="cuda", dtype=torch.bool) for _ in range(2)]
        assert pair.run(sb, ob, L.ANY) == [0, 0], L.lib().b200phi_last_error()
        assert np.array_equal(G.to_np(ob[0]), fb.any(0)) and torch.equal(ob[0], ob[1])
        assert pair.run(sb, ob, L.ALL) == [0, 0]
        assert np.array_equal(G.to_np(ob[1]), fb.all(0))
    finally:
        pair.close()


_TIMEOUT_SCRIPT = r"""
import ctypes as C, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {tests!r})
import torch
from paddle2_b200 import _lib as L
import test_p2p_one_gpu as T
pair = T._Pair(max_out=512)
x = [torch.rand(5000, 512, device="cuda"), torch.rand(5000, 512, device="cuda")]
outs = [torch.full((512,), float("nan"), device="cuda") for _ in range(2)]
rcs = pair.run(x, outs, L.SUM, only_rank=0)            # rank 1 never calls
assert rcs == [0], rcs
assert bool(torch.isnan(outs[0]).all()), "nothing may be combined after a timeout"
assert L.lib().b200phi_p2p_status(pair.bufs[0]) == 6, "status must report PEER_TIMEOUT"
assert b"did not publish" in L.lib().b200phi_last_error()
rcs = pair.run(x, outs, L.SUM, only_rank=0)            # the next call on the group refuses
assert rcs == [6], rcs
assert L.lib().b200phi_p2p_status(pair.bufs[1]) == 0
pair.close()
print("timeout ok")
"""


def test_a_silent_peer_is_a_timeout_not_a_wrong_sum():
    env = dict(os.environ, B200PHI_P2P_TIMEOUT_MS="150")
    script = _TIMEOUT_SCRIPT.format(root=ROOT, tests=os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", script], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "timeout ok" in r.stdout, r.stdout + r.stderr
