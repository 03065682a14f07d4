"""Host-buffer entry point: reduce a tensor that lives in (pinned) HOST memory.

This is the call bench.py times end to end: the rows stream over PCIe in chunks on a
copy stream while the previous chunk is being reduced on the compute stream
(b200phi_reduce per chunk into one row of a [chunks, C] partial matrix), and a last
b200phi_reduce folds the partial matrix.  Only the [C] result travels back.
torch provides pinned memory, streams and events; the arithmetic is libb200phi.so's.
"""
import torch

from . import ops


class HostAxis0Reducer:
    """sum/max/min/... over axis 0 of a [rows, cols] host tensor, double-buffered."""

    def __init__(self, cols, dtype=torch.float32, device="cuda", chunk_rows=16384, op="sum"):
        self.device = torch.device(device)
        self.cols = cols
        self.dtype = dtype
        self.chunk_rows = chunk_rows
        self.op = getattr(ops, op)
        self.buf = [torch.empty(chunk_rows, cols, dtype=dtype, device=self.device) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.free_evt = [torch.cuda.Event() for _ in range(2)]
        self.full_evt = [torch.cuda.Event() for _ in range(2)]
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _out_dtype(self):
        if self.op is ops.sum:
            return ops._sum_out_dtype(self.dtype, None)
        return self.dtype

    def __call__(self, x_host, out_host=None):
        if x_host.is_cuda:
            raise ValueError("HostAxis0Reducer takes a host tensor; use ops.sum for device tensors")
        if x_host.dim() != 2 or x_host.shape[1] != self.cols or x_host.dtype != self.dtype:
            raise ValueError("host tensor must be [rows, cols] of the configured dtype")
        rows = x_host.shape[0]
        if rows == 0:
            raise ValueError("Tensor need be reduced must not empty.")
        nchunks = (rows + self.chunk_rows - 1) // self.chunk_rows
        compute = torch.cuda.current_stream(self.device)
        parts = torch.empty(nchunks, self.cols, dtype=self._out_dtype(), device=self.device)
        self.h2d_bytes = self.d2h_bytes = 0
        for i in range(nchunks):
            r0 = i * self.chunk_rows
            r1 = min(r0 + self.chunk_rows, rows)
            b = i & 1
            with torch.cuda.stream(self.copy_stream):
                if i >= 2:
                    self.copy_stream.wait_event(self.free_evt[b])   # chunk i-2 has been reduced
                dst = self.buf[b][: r1 - r0]
                dst.copy_(x_host[r0:r1], non_blocking=True)
                self.full_evt[b].record(self.copy_stream)
            self.h2d_bytes += (r1 - r0) * self.cols * x_host.element_size()
            compute.wait_event(self.full_evt[b])
            self.op(self.buf[b][: r1 - r0], axis=0, out=parts[i])
            self.free_evt[b].record(compute)
        res = parts[0] if nchunks == 1 else self.op(parts, axis=0)
        if out_host is None:
            out_host = torch.empty(res.shape, dtype=res.dtype, pin_memory=True)
        out_host.copy_(res, non_blocking=True)
        self.d2h_bytes = res.numel() * res.element_size()
        compute.synchronize()
        return out_host
