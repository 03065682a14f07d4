"""One launch of each short-row / flat / moments kernel, for an ncu capture."""
import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import ops
x = torch.randn(1 << 20, 256, device="cuda", dtype=torch.bfloat16)
pb.max(x, axis=-1); pb.row_mean_var(x); ops.logsumexp(x, -1); pb.argmax(x, axis=-1)
del x
x = torch.randn(16384, 8192, device="cuda", dtype=torch.bfloat16)
pb.row_mean_var(x)
del x
x = torch.randn(1 << 21, 64, device="cuda"); y = torch.randn(64, device="cuda")
pb.add(x, y); pb.less_than(x, y)
del x
x = torch.randn(10699, 64, 196, device="cuda"); y = torch.randn(64, 1, device="cuda")
pb.fused_elemwise_add_activation(x, y, ["relu", "elementwise_add"])
xb = x.bfloat16(); yb = y.bfloat16()
pb.add(xb, yb)
del x, xb
x = torch.randn(64, 64, 112, 112, device="cuda")
pb.sum(x, axis=[0, 2, 3])
del x
x = torch.randn(8192, 50257, device="cuda", dtype=torch.bfloat16)
ops.logsumexp(x, -1); pb.row_mean_var(x)
torch.cuda.synchronize()
print("ok")
