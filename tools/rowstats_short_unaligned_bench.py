import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
def t(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
N = 1 << 30
for dt in (torch.float32, torch.bfloat16):
    for cols in (201, 301, 501, 701, 1001):
        rows = N // cols
        x = torch.randn(rows, cols, device="cuda").to(dt)
        us = t(lambda: pb.row_mean_var(x), 3)
        print(f"mean+var {str(dt)[6:]:9s} [{rows},{cols}] {cols*x.element_size():6d} B/row {us:9.1f} us {x.numel()*x.element_size()/us/1e6:6.2f} TB/s", flush=True)
        del x
