import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
N = 1 << 29
for cols in (16, 96, 128):
    x = torch.randn(N // cols, cols, device="cuda").bfloat16()
    pb.sum(x, axis=-1)
    del x
torch.cuda.synchronize()
