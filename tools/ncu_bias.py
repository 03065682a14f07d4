import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import grad
for dt in (torch.bfloat16, torch.float32):
    x = torch.randn(65536, 4096, device="cuda").to(dt)
    y = torch.randn(4096, device="cuda").to(dt)
    pb.add(x, y)
    out = pb.max(x, axis=0, keepdim=True)
    g = torch.randn(1, 4096, device="cuda").to(dt)
    grad.max_grad(x, out, g, [0], True)
    del x
torch.cuda.synchronize()
