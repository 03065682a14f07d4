import torch, sys
sys.path.insert(0, '.')
import paddle2_b200 as pb
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for dt, rows, cols in [(torch.bfloat16, 524288, 784), (torch.float32, 262144, 784), (torch.bfloat16, 131072, 3136), (torch.float32, 65536, 3136), (torch.bfloat16, 1048576, 392), (torch.float32, 1048576, 196), (torch.float32, 16384, 12544)]:
    x = torch.randn(rows, cols, device='cuda').to(dt)
    out = torch.empty(rows, device='cuda', dtype=torch.float32 if dt != torch.float32 else dt)
    us = t(lambda: pb.sum(x, axis=-1, dtype='float32', out=out))
    print(f"{dt} [{rows},{cols}] row sum: {us:8.1f} us  {x.numel()*x.element_size()/us/1e6:8.1f} TB/s")
