"""Sweep of mid-sized shapes through the public ops (GPU time by CUDA events, tensors of
~0.5 GiB so the kernel dwarfs the ctypes call): finds the shapes between the tuned configs."""
import sys
import torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import ops


def t(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


def line(name, us, nbytes):
    print(f"{name:58s} {us:9.1f} us {nbytes / us / 1e6:7.2f} TB/s", flush=True)


which = sys.argv[1] if len(sys.argv) > 1 else "all"
N = 1 << 28  # elements for 16-bit, halves for fp32
if which in ("all", "ew"):
    for hw in (49, 196, 784, 3136, 12544):
        c = 64
        n = (N // 2) // (c * hw)
        x = torch.randn(n, c, hw, device="cuda")
        y = torch.randn(c, 1, device="cuda")
        line(f"relu(x+y) fp32 [{n},{c},{hw}] + [{c},1]", t(lambda: pb.fused_elemwise_add_activation(x, y, ["relu", "elementwise_add"])), x.numel() * 8)
        xb = x.bfloat16(); yb = y.bfloat16()
        line(f"x+y bf16 [{n},{c},{hw}] + [{c},1]", t(lambda: pb.add(xb, yb)), xb.numel() * 4)
        del x, xb
    for inner in (64, 768, 4096):
        rows = (N // 2) // inner
        x = torch.randn(rows, inner, device="cuda"); y = torch.randn(inner, device="cuda")
        line(f"x+y fp32 [{rows},{inner}] + [{inner}]", t(lambda: pb.add(x, y)), x.numel() * 8)
        line(f"less_than fp32 [{rows},{inner}] < [{inner}]", t(lambda: pb.less_than(x, y)), x.numel() * 5)
        del x
if which in ("all", "stats"):
    for cols in (256, 768, 1024, 2048, 4096, 8192):
        rows = N // cols
        x = torch.randn(rows, cols, device="cuda").bfloat16()
        line(f"row mean+var bf16 [{rows},{cols}]", t(lambda: pb.row_mean_var(x)), x.numel() * 2)
        line(f"logsumexp bf16 [{rows},{cols}]", t(lambda: ops.logsumexp(x, -1)), x.numel() * 2)
        line(f"argmax bf16 [{rows},{cols}]", t(lambda: pb.argmax(x, axis=-1)), x.numel() * 2)
        line(f"max bf16 [{rows},{cols}]", t(lambda: pb.max(x, axis=-1)), x.numel() * 2)
        del x
if which in ("all", "col"):
    for r, c in ((64, 1 << 21), (1024, 1 << 17), (16384, 8192), (1 << 20, 128), (1 << 22, 32)):
        x = torch.randn(r, c, device="cuda")
        line(f"sum fp32 [{r},{c}] axis 0", t(lambda: pb.sum(x, axis=0)), x.numel() * 4)
        del x
    for a, k, b in ((32, 1024, 4096), (8, 64, 262144), (1024, 32, 4096)):
        x = torch.randn(a, k, b, device="cuda")
        line(f"sum fp32 [{a},{k},{b}] axis 1", t(lambda: pb.sum(x, axis=1)), x.numel() * 4)
        del x
