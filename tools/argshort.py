"""argmax over aligned 16-bit rows of 64..2048 columns (2 GiB tensors)."""
import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
def t(fn, n=3):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
N = 1 << 30
for dt in (torch.bfloat16, torch.float16):
    for cols in (64, 128, 256, 384, 512, 768, 1024, 1280, 1536, 2048):
        x = torch.randn(N // cols, cols, device="cuda").to(dt)
        us = t(lambda: pb.argmax(x, axis=-1))
        us2 = t(lambda: pb.argmax_with_value(x, axis=-1))
        print(f"argmax {str(dt)[6:]:9s} [{N//cols},{cols}] {x.numel()*2/us/1e6:5.2f} TB/s   with value {x.numel()*2/us2/1e6:5.2f}", flush=True)
        del x
