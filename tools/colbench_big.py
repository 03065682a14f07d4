"""Column / middle-axis reductions on 4 GiB tensors (kernel-bound, not call-bound)."""
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
for shape, ax in (((512, 1 << 21), 0), ((8192, 1 << 17), 0), ((131072, 8192), 0), ((1 << 23, 128), 0), ((1 << 25, 32), 0),
                  ((256, 1024, 4096), 1), ((64, 64, 262144), 1), ((8192, 32, 4096), 1), ((256, 64, 112 * 112 * 4), (0, 2))):
    x = torch.randn(shape, device="cuda")
    us = t(lambda: pb.sum(x, axis=ax))
    print(f"sum fp32 {list(shape)} axis {ax}: {us:9.1f} us {x.numel()*4/us/1e6:6.2f} TB/s", flush=True)
    del x
