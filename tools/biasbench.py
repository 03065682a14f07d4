"""Row-vector broadcasts (bias add and friends): [N, C] op [C], and max_grad along axis 0 where two
[C] operands ride along; 1-4 GiB tensors, CUDA events over 3 calls."""
import sys
import torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import grad


def t(fn, n=3):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


for dt in (torch.float32, torch.bfloat16):
    es = torch.empty((), dtype=dt).element_size()
    for rows, cols in ((262144, 1024), (65536, 4096), (16384, 16384), (4096, 65536)):
        x = torch.randn(rows, cols, device="cuda").to(dt)
        y = torch.randn(cols, device="cuda").to(dt)
        us = t(lambda: pb.add(x, y))
        line = f"{str(dt)[6:]:9s} [{rows},{cols}] + [{cols}]  add {2 * x.numel() * es / us / 1e6:5.2f}"
        us = t(lambda: pb.fused_elemwise_add_activation(x, y, ["relu", "elementwise_add"]))
        line += f"  relu(x+y) {2 * x.numel() * es / us / 1e6:5.2f}"
        us = t(lambda: pb.less_than(x, y))
        line += f"  less_than {(x.numel() * es + x.numel()) / us / 1e6:5.2f}"
        out = pb.max(x, axis=0, keepdim=True)
        g = torch.randn(1, cols, device="cuda").to(dt)
        us = t(lambda: grad.max_grad(x, out, g, [0], True))
        line += f"  max_grad axis0 {2 * x.numel() * es / us / 1e6:5.2f}"
        print(line + "  TB/s", flush=True)
        del x
