"""Same-shape elementwise and unary maps, fp32 and bf16, 1 GiB tensors."""
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
for dt in (torch.float32, torch.bfloat16):
    es = 4 if dt == torch.float32 else 2
    x = torch.randn((1 << 30) // es, device="cuda").to(dt).view(-1, 4096)
    y = torch.randn_like(x)
    line = f"{str(dt)[6:]:9s} [{x.shape[0]},4096]"
    for name, fn, nb in (("add", lambda: pb.add(x, y), 3), ("multiply", lambda: pb.multiply(x, y), 3),
                         ("relu", lambda: pb.relu(x), 2), ("exp", lambda: pb.exp(x), 2), ("silu", lambda: pb.silu(x), 2),
                         ("tanh", lambda: pb.tanh(x), 2), ("sigmoid", lambda: pb.sigmoid(x), 2),
                         ("less_than", lambda: pb.less_than(x, y), 2 + 1.0 / es)):
        us = t(fn)
        line += f"  {name} {nb * x.numel() * es / us / 1e6:5.2f}"
    print(line + "  TB/s", flush=True)
