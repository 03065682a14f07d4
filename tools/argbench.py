"""argmax shapes where the reduced extent is split across CTAs: flatten of a long
tensor, axis 0 of a row shard, a middle axis.  CUDA events over `iters` back-to-back calls through
the C ABI (descriptor and workspace prepared once).  GB/s = input bytes / time."""
import ctypes as C
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import paddle2_b200 as pb
from paddle2_b200 import _lib as L


def bench(x, axis, iters=10, with_value=False):
    lib = L.lib()
    d = L.ArgReduceDesc()
    d.op, d.in_dtype, d.index_dtype, d.rank = L.ARGMAX, pb.ops._phi_dtype(x.dtype), L.INT64, x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.flatten = 1 if axis is None else 0
    d.axis = 0 if axis is None else axis
    oshape = [] if axis is None else [s for i, s in enumerate(x.shape) if i != axis]
    out = torch.empty(oshape, device="cuda", dtype=torch.int64)
    val = torch.empty(oshape, device="cuda", dtype=x.dtype)
    d.x, d.out = x.data_ptr(), out.data_ptr()
    d.max_out = val.data_ptr() if with_value else None
    ws_bytes = lib.b200phi_argreduce_workspace(C.byref(d))
    ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device="cuda")
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for _ in range(3):
        L.check(lib.b200phi_argreduce(C.byref(d), st))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        L.check(lib.b200phi_argreduce(C.byref(d), st))
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    gbs = x.numel() * x.element_size() / us / 1e3
    return us, gbs, ws_bytes, out


def main():
    rows = []
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(1 << 30, device="cuda", generator=g).bfloat16()
    us, gbs, ws, out = bench(x, None)
    assert int(out) == int(torch.argmax(x.float()))
    rows.append(("argmax flatten bf16 2^30", us, gbs, ws))
    del x
    x = torch.randint(-(2 ** 40), 2 ** 40, (1 << 29,), device="cuda", dtype=torch.int64)
    us, gbs, ws, out = bench(x, None)
    assert int(out) == int(torch.argmax(x))
    rows.append(("argmax flatten int64 2^29", us, gbs, ws))
    del x
    x = torch.randn(524288, 4096, device="cuda", generator=g)
    us, gbs, ws, out = bench(x, 0, iters=5, with_value=True)
    assert torch.equal(x.gather(0, out[None, :])[0], x.max(0).values)
    rows.append(("argmax+max fp32 [524288,4096] axis 0", us, gbs, ws))
    del x
    x = torch.randn(262144, 8192, device="cuda", generator=g).bfloat16()
    us, gbs, ws, out = bench(x, 0, iters=5)
    rows.append(("argmax bf16 [262144,8192] axis 0", us, gbs, ws))
    del x
    x = torch.randn(64, 21, 512 * 512, device="cuda", generator=g)
    us, gbs, ws, out = bench(x, 1)
    assert torch.equal(out, x.argmax(1))
    rows.append(("argmax fp32 [64,21,512*512] axis 1 (segmentation classes)", us, gbs, ws))
    del x
    x = torch.randn(256, 4096, 1024, device="cuda", generator=g).bfloat16()
    us, gbs, ws, out = bench(x, 1)
    rows.append(("argmax bf16 [256,4096,1024] axis 1", us, gbs, ws))
    for name, us, gbs, ws in rows:
        print(f"{name:58s} {us:10.1f} us {gbs:8.1f} GB/s  ws {ws} B")


if __name__ == "__main__":
    main()
