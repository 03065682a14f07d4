"""Probe: the fused-exchange kernel with a one-rank group against the plain column reduce
on the C5 shape (isolates the cost of the epilogue from cross-GPU skew)."""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle2_b200 import _lib as L  # noqa: E402
from paddle2_b200.dist import P2PReducer  # noqa: E402

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29544")
dist.init_process_group("gloo", rank=0, world_size=1)
R, Cc = 524288, 4096
x = torch.rand(R, Cc, device="cuda")
out = torch.empty(Cc, device="cuda")
lib = L.lib()
d = L.ReduceDesc()
d.op, d.in_dtype, d.out_dtype, d.rank = L.SUM, L.FLOAT32, L.FLOAT32, 2
d.dims[0], d.dims[1] = R, Cc
d.naxes = 1
d.axes[0] = 0
d.x, d.out = x.data_ptr(), out.data_ptr()
ws = torch.empty(lib.b200phi_reduce_workspace(C.byref(d)), dtype=torch.uint8, device="cuda")
d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
p2p = P2PReducer(max_out_numel=Cc, max_tiles=64)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def plain():
    L.check(lib.b200phi_reduce(C.byref(d), st))


def fused():
    p2p.epoch += 1
    L.check(lib.b200phi_sharded_reduce_p2p(C.byref(d), 0, C.byref(p2p.g), p2p.epoch, st))


for name, fn in (("plain", plain), ("fused(1 rank)", fused), ("plain", plain), ("fused(1 rank)", fused)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:15s} {e0.elapsed_time(e1) / 20 * 1e3:9.1f} us")
ref = x.double().sum(0).float()
fused()
torch.cuda.synchronize()
print("max rel err", float(((out - ref).abs() / ref).max()))
p2p.close()
dist.destroy_process_group()
