"""mean+var and logsumexp over aligned 16-bit rows of 256..1024 columns (2 GiB tensors)."""
import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import ops
def t(fn, n=3):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
N = 1 << 30
for cols in (136, 200, 256, 384, 512, 768, 1024, 1280, 1536, 2048):
    x = torch.randn(N // cols, cols, device="cuda").bfloat16()
    us = t(lambda: pb.row_mean_var(x))
    us2 = t(lambda: ops.logsumexp(x, -1))
    print(f"bf16 [{N//cols},{cols}]  mean+var {x.numel()*2/us/1e6:5.2f}  logsumexp {x.numel()*2/us2/1e6:5.2f} TB/s", flush=True)
    del x
