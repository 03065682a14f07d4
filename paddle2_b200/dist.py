"""Sharded global reduce over the GPUs of one box: one process per GPU.

Config C5: a tensor too large for one GPU is split along axis 0 into contiguous
row blocks; each rank reduces its block locally and the [C]-sized partials are
combined by an allreduce over NVLink issued on the same CUDA stream.  This is the
reference's `paddle.sum(local, axis=0)` + `paddle.distributed.all_reduce`
(process_group_nccl.cc:267-298) / Partial->Replicate reshard
(p_to_r_reshard_function.cc:67) as ONE C-ABI call, b200phi_sharded_reduce.

torch.distributed is only the rendezvous: it carries the 128-byte NCCL unique id
from rank 0 to the others (what CommContextManager does over its TCPStore,
comm_context_manager.cc:75) and the row counts.  The data path never goes through
torch collectives.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib as L
from . import ops as _ops


def partition_rows(total_rows, world_size, rank):
    """Contiguous row block of `rank`: (start, count).  Blocks differ by at most one
    row and cover [0, total_rows) exactly (the layout Shard(0) gives a DistTensor)."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank/world_size")
    base, extra = divmod(total_rows, world_size)
    start = rank * base + (rank if rank < extra else extra)
    return start, base + (1 if rank < extra else 0)


def exchange_unique_id(make_id, group=None):
    """Rank 0 calls make_id() -> bytes(128); every rank returns those bytes."""
    rank = dist.get_rank(group)
    box = [make_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    uid = box[0]
    if not isinstance(uid, (bytes, bytearray)) or len(uid) != 128:
        raise RuntimeError("unique id rendezvous failed")
    return bytes(uid)


def global_count(local_rows, group=None):
    """Sum of the ranks' row counts (the divisor of a sharded mean)."""
    t = torch.tensor([int(local_rows)], dtype=torch.int64)
    backend = dist.get_backend(group)
    if backend == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return int(t.item())


def row_offset_of(local_rows, group=None):
    """Global index of this rank's first row: exclusive prefix sum of every rank's row count
    (host-side, over the rendezvous group; works with gloo and nccl alike)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return 0
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    counts = [None] * world
    dist.all_gather_object(counts, int(local_rows), group=group)
    off = 0
    for r in range(rank):
        off += counts[r]
    return off


class ShardedReducer:
    """Owns one NCCL communicator created through the C ABI (b200phi_comm_*)."""

    def __init__(self, group=None):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed must be initialised (it is the rendezvous)")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        lib = L.lib()
        self.comm = None
        if self.world == 1:
            return  # single GPU: no communicator, no collective (north star)

        def make_id():
            buf = C.create_string_buffer(128)
            L.check(lib.b200phi_comm_unique_id(buf))
            return buf.raw

        uid = exchange_unique_id(make_id, group)
        comm = C.c_void_p()
        L.check(lib.b200phi_comm_init_rank(uid, self.world, self.rank, C.byref(comm)))
        self.comm = comm

    def close(self):
        if getattr(self, "comm", None):
            L.check(L.lib().b200phi_comm_destroy(self.comm))
            self.comm = None

    def reduce(self, op, x_local, global_rows=None, out=None, workspace=None):
        """Reduce the row-sharded tensor over axis 0; every rank gets the result.
        `op` is a b200phi reduce op id (L.SUM, L.MEAN, L.MAX, ...)."""
        _ops._require_cuda(x_local)
        x = x_local.contiguous()
        if x.dim() < 1:
            raise ValueError("sharded reduce needs rank >= 1")
        out_dtype = _ops._sum_out_dtype(x.dtype, None) if op == L.SUM else x.dtype
        if op in (L.ANY, L.ALL):
            out_dtype = torch.bool
        if out is None:
            out = torch.empty(x.shape[1:], dtype=out_dtype, device=x.device)
        d = L.ReduceDesc()
        d.op = op
        d.in_dtype = _ops._phi_dtype(x.dtype)
        d.out_dtype = _ops._phi_dtype(out.dtype)
        d.rank = x.dim()
        for i, s in enumerate(x.shape):
            d.dims[i] = s
        d.naxes = 1
        d.axes[0] = 0
        d.reduce_all = 0
        d.x = x.data_ptr()
        d.out = out.data_ptr()
        lib = L.lib()
        with torch.cuda.device(x.device):
            ws_bytes = lib.b200phi_sharded_reduce_workspace(C.byref(d))
            if ws_bytes:
                if workspace is None or workspace.numel() < ws_bytes:
                    workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
                d.workspace = workspace.data_ptr()
                d.workspace_bytes = workspace.numel()
            if global_rows is None:
                global_rows = global_count(x.shape[0], self.group) if op == L.MEAN else 0
            stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            L.check(lib.b200phi_sharded_reduce(C.byref(d), int(global_rows), self.comm, stream))
        return out

    def sum(self, x_local, **kw):  # noqa: A003
        return self.reduce(L.SUM, x_local, **kw)

    def argreduce(self, op, x_local, row_offset=None, dtype="int64", want_value=False):
        """argmax / argmin over axis 0 of the row-sharded tensor; every rank gets the global row
        index per column (lowest index on ties) and, optionally, the extreme value.
        `row_offset`: global index of this shard's first row; by default the exclusive prefix sum
        of the shards' row counts, exchanged once over the rendezvous group."""
        _ops._require_cuda(x_local)
        x = x_local.contiguous()
        if x.dim() < 1:
            raise ValueError("sharded argreduce needs rank >= 1")
        if row_offset is None:
            row_offset = row_offset_of(x.shape[0], self.group)
        idx_dtype = _ops._to_torch_dtype(dtype)
        out = torch.empty(x.shape[1:], dtype=idx_dtype, device=x.device)
        val = torch.empty(x.shape[1:], dtype=x.dtype, device=x.device) if want_value else None
        d = L.ArgReduceDesc()
        d.op = op
        d.in_dtype = _ops._phi_dtype(x.dtype)
        d.index_dtype = _ops._phi_dtype(idx_dtype)
        d.rank = x.dim()
        for i, s in enumerate(x.shape):
            d.dims[i] = s
        d.axis = 0
        d.flatten = 0
        d.x = x.data_ptr()
        d.out = out.data_ptr()
        d.max_out = val.data_ptr() if want_value else None
        lib = L.lib()
        with torch.cuda.device(x.device):
            ws_bytes = lib.b200phi_sharded_argreduce_workspace(C.byref(d), self.world)
            ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device=x.device)
            stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            L.check(lib.b200phi_sharded_argreduce(C.byref(d), int(row_offset), self.world, self.comm,
                                                  ws.data_ptr(), ws.numel(), stream))
        return (out, val) if want_value else out

    def argmax(self, x_local, **kw):
        return self.argreduce(L.ARGMAX, x_local, **kw)

    def argmin(self, x_local, **kw):
        return self.argreduce(L.ARGMIN, x_local, **kw)


class P2PReducer:
    """Sharded reduce without a collective library: local reduce + a small exchange kernel over
    NVLink peer memory (b200phi_sharded_reduce_p2p).  torch.distributed only carries the 64-byte
    CUDA IPC handles at set-up; the data path is the kernel's own peer stores and loads."""

    def __init__(self, max_out_numel, max_tiles=64, group=None):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed must be initialised (it is the rendezvous)")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if self.world > 8:
            raise ValueError("the fused exchange covers one NVSwitch box: at most 8 ranks")
        lib = L.lib()
        self.lib = lib
        self.max_out, self.max_tiles = int(max_out_numel), int(max_tiles)
        nbytes = lib.b200phi_p2p_buffer_bytes(self.world, self.max_out, self.max_tiles)
        own = C.c_void_p()
        handle = C.create_string_buffer(64)
        L.check(lib.b200phi_p2p_alloc(nbytes, C.byref(own), handle))
        self.own = own
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw, group=group)
        self.peers = []
        g = L.P2PGroup()
        g.nranks, g.rank = self.world, self.rank
        g.max_out_numel, g.max_tiles = self.max_out, self.max_tiles
        for r, h in enumerate(handles):
            if r == self.rank:
                g.base[r] = own.value
            else:
                p = C.c_void_p()
                L.check(lib.b200phi_p2p_open(h, C.byref(p)))
                self.peers.append(p)
                g.base[r] = p.value
        self.g = g
        self.epoch = 0
        dist.barrier(group=group)  # every rank has mapped every buffer before the first kernel

    def close(self):
        if getattr(self, "own", None) is None:
            return
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        for p in self.peers:
            self.lib.b200phi_p2p_close(p)
        self.lib.b200phi_p2p_free(self.own)
        self.own, self.peers = None, []

    def status(self):
        """Raises B200PhiError(PEER_TIMEOUT) if an exchange gave up waiting for a peer."""
        L.check(self.lib.b200phi_p2p_status(self.own))

    def reduce(self, op, x_local, global_rows=0, out=None, workspace=None):
        _ops._require_cuda(x_local)
        x = x_local.contiguous()
        out_dtype = _ops._sum_out_dtype(x.dtype, None) if op == L.SUM else x.dtype
        if op in (L.ANY, L.ALL):
            out_dtype = torch.bool
        if out is None:
            out = torch.empty(x.shape[1:], dtype=out_dtype, device=x.device)
        d = L.ReduceDesc()
        d.op = op
        d.in_dtype = _ops._phi_dtype(x.dtype)
        d.out_dtype = _ops._phi_dtype(out.dtype)
        d.rank = x.dim()
        for i, s in enumerate(x.shape):
            d.dims[i] = s
        d.naxes = 1
        d.axes[0] = 0
        d.x = x.data_ptr()
        d.out = out.data_ptr()
        with torch.cuda.device(x.device):
            ws_bytes = self.lib.b200phi_reduce_workspace(C.byref(d))
            if ws_bytes:
                if workspace is None or workspace.numel() < ws_bytes:
                    workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
                d.workspace = workspace.data_ptr()
                d.workspace_bytes = workspace.numel()
            self.epoch += 1
            stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            L.check(self.lib.b200phi_sharded_reduce_p2p(C.byref(d), int(global_rows), C.byref(self.g),
                                                        self.epoch, stream))
        return out

    def sum(self, x_local, **kw):  # noqa: A003
        return self.reduce(L.SUM, x_local, **kw)
