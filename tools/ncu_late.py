"""One launch of each kernel added late in round 2, for an ncu capture:
  ncu --set full --clock-control none -k regex:'rows_staged|flat_row|warp_peel|row_peel' -o /tmp/late python tools/ncu_late.py
1 GiB inputs."""
import sys
import torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import grad, ops

G = 1 << 30
x = torch.randn(G // 4 // 32, 32, device="cuda")
pb.row_mean_var(x)                      # rows_staged_kernel<float, StagedStats>, aligned rows (LDS.128, staggered)
ops.logsumexp(x, -1)                    # rows_staged_kernel<float, StagedLse>
del x
x = torch.randn(G // 2 // 64, 64, device="cuda").bfloat16()
pb.argmax(x, axis=-1)                   # rows_staged_kernel<bf16, StagedArg>, aligned
del x
x = torch.randn(G // 4 // 8, 8, device="cuda")
pb.argmax(x, axis=-1)                   # rows_staged_kernel, 8 rows per thread and span
del x
x = torch.randn(65536, 8192, device="cuda").bfloat16()
y = torch.randn(8192, device="cuda").bfloat16()
pb.add(x, y)                            # bcast_flat_row_kernel (bias add), wait: 1024 vectors per row -> rows kernel
x2 = x[:, :4096].contiguous()
y2 = y[:4096].contiguous()
pb.add(x2, y2)                          # bcast_flat_row_kernel
pb.less_than(x2, y2)                    # map_flat_row_kernel (binary)
out = pb.max(x2, axis=0, keepdim=True)
g = torch.randn(1, 4096, device="cuda").bfloat16()
grad.max_grad(x2, out, g, [0], True)    # map_flat_row_kernel (two shared rows)
del x, x2
x = torch.randn(G // 2 // 2048, 2048, device="cuda").bfloat16()
pb.argmax(x, axis=-1)                   # arg_warp_peel_kernel<..., 8> on aligned rows
del x
x = torch.randn(G // 2 // 4097 * 4097 + 1, device="cuda").bfloat16()[1:].view(-1, 4097)
pb.sum(x, axis=-1)                      # row_peel_kernel<..., 32, 16>
torch.cuda.synchronize()
print("ok")
