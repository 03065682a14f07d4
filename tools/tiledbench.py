"""General keep/reduce patterns: GB/s of b200phi_reduce through the C ABI, CUDA events
over `iters` back-to-back calls, descriptor and workspace prepared once.  Prints the plan kind
(10 = tile_keep_kernel, 11 = tile_red_kernel, 3 = col_reduce_kernel, 6 = two-stage)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import paddle2_b200 as pb
from paddle2_b200 import _lib as L


def bench(x, axes, op=L.SUM, iters=10):
    lib = L.lib()
    d = L.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = op, pb.ops._phi_dtype(x.dtype), 0, x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.naxes = len(axes)
    for i, a in enumerate(axes):
        d.axes[i] = a
    oshape = [s for i, s in enumerate(x.shape) if i not in axes]
    out = torch.empty(oshape, device="cuda", dtype=x.dtype)
    d.x, d.out = x.data_ptr(), out.data_ptr()
    kind = lib.b200phi_reduce_plan_kind(C.byref(d))
    ws_bytes = lib.b200phi_reduce_workspace(C.byref(d))
    ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device="cuda")
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for _ in range(3):
        L.check(lib.b200phi_reduce(C.byref(d), st))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        L.check(lib.b200phi_reduce(C.byref(d), st))
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    nbytes = x.numel() * x.element_size() + out.numel() * out.element_size()
    return us, nbytes / us / 1e3, kind, ws_bytes, out


def main():
    g = torch.Generator(device="cuda").manual_seed(1)
    cases = [
        ("sum fp32 [256,64,112,112] axes (0,2)", (256, 64, 112, 112), [0, 2], torch.float32, L.SUM),
        ("sum fp32 [256,64,112,112] axes (1,3)", (256, 64, 112, 112), [1, 3], torch.float32, L.SUM),
        ("min fp32 [32,64,48,80,6,10] axes (2,4)", (32, 64, 48, 80, 6, 10), [2, 4], torch.float32, L.MIN),
        ("prod fp32 [40,10,6,7,8,9,10,22] axes (2,3,4)", (40, 10, 6, 7, 8, 9, 10, 22), [2, 3, 4], torch.float32, L.PROD),
        ("sum bf16 [256,64,112,112] axes (0,2)", (256, 64, 112, 112), [0, 2], torch.bfloat16, L.SUM),
        ("sum bf16 [256,64,112,112] axes (1,3)", (256, 64, 112, 112), [1, 3], torch.bfloat16, L.SUM),
        ("sum fp32 [8, 16M, 8] axis 0 (R<16 column)", (8, 1 << 24, 8), [0], torch.float32, L.SUM),
        ("sum fp32 [4M, 16, 8] axes (0,2) (short inner run)", (1 << 22, 16, 8), [0, 2], torch.float32, L.SUM),
        ("sum fp32 [64,12,4096,64] axes (1,3) (heads x dim)", (64, 12, 4096, 64), [1, 3], torch.float32, L.SUM),
        ("max fp32 [512,7,7,2048] axes (1,2) (NHWC pool)", (512, 7, 7, 2048), [1, 2], torch.float32, L.MAX),
    ]
    for name, shape, axes, dt, op in cases:
        if op == L.PROD:
            x = (0.95 + 0.1 * torch.rand(shape, device="cuda", generator=g)).to(dt)
        else:
            x = torch.rand(shape, device="cuda", generator=g).to(dt)
        us, gbs, kind, ws, out = bench(x, axes, op)
        if op == L.SUM:
            want = x.double().sum(dim=axes)
            tol = 1e-5 if dt == torch.float32 else 1e-2
            assert bool(((out.double() - want).abs() <= tol * want.abs()).all()), name
        elif op == L.MIN:
            assert torch.equal(out, x.amin(dim=axes)), name
        elif op == L.MAX:
            assert torch.equal(out, x.amax(dim=axes)), name
        print(f"{name:56s} kind {kind:2d} {us:9.1f} us {gbs:8.1f} GB/s  ws {ws} B  ({x.numel() * x.element_size() / 2**20:.0f} MiB)")
        del x, out
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
