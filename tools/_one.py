import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
x = torch.randn(174762, 768, device="cuda"); y = torch.randn(768, device="cuda")
for _ in range(2):
    o = pb.less_than(x, y)
torch.cuda.synchronize()
x2 = torch.randn(2097152, 64, device="cuda"); y2 = torch.randn(64, device="cuda")
o = pb.less_than(x2, y2)
torch.cuda.synchronize()
print("ok")
