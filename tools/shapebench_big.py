import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import ops
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
N = 1 << 31
for cols in (256, 512, 768, 1024, 2048, 4096):
    rows = N // cols
    x = torch.randn(rows, cols, device="cuda", dtype=torch.bfloat16)
    for name, fn in (("mean+var", lambda: pb.row_mean_var(x)), ("logsumexp", lambda: ops.logsumexp(x, -1)),
                     ("argmax", lambda: pb.argmax(x, axis=-1)), ("max", lambda: pb.max(x, axis=-1))):
        us = t(fn, 5)
        print(f"{name:10s} bf16 [{rows},{cols}] {us:9.1f} us {x.numel()*2/us/1e6:6.2f} TB/s", flush=True)
    del x
