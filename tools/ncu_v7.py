"""One launch of each kernel that serves the shapes VERDICT r1 #7 names (16-bit rows that are not
whole vectors, unaligned 16-32 KiB statistics rows, ragged broadcasts against a per-row scalar),
for an ncu capture:
  ncu --set full --clock-control none -k regex:'peel|scalar_kernel|rowstats_reg|ragged_rowscalar|rows_staged' \
      -o /tmp/v7 python tools/ncu_v7.py && python tools/ncu_summary.py /tmp/v7.ncu-rep
~1 GiB tensors (ncu replays each kernel ~40 times)."""
import sys
import torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import ops

N = 1 << 29
x = torch.randn(N // 501, 501, device="cuda").bfloat16()
pb.sum(x, axis=-1)            # row_peel_kernel
pb.argmax(x, axis=-1)         # arg_warp_peel_kernel (sub-warp groups)
pb.row_mean_var(x)            # rowstats_warp_peel_kernel (sub-warp groups)
ops.logsumexp(x, -1)          # lse_peel_kernel
del x
x = torch.randn(N // 101, 101, device="cuda").bfloat16()   # 202-byte rows: 12 whole vectors at most
pb.sum(x, axis=-1); pb.max(x, axis=-1); pb.argmax(x, axis=-1); pb.row_mean_var(x); ops.logsumexp(x, -1)
del x
x = torch.randn(N // 10001, 10001, device="cuda").bfloat16()
pb.row_mean_var(x)            # rowstats_reg_kernel<..., kPeel>
del x
x = torch.randn(N // 2 // 5001, 5001, device="cuda")
pb.row_mean_var(x)            # rowstats_reg_kernel<float, ..., kPeel>
del x
n = N // (64 * 49)
x = torch.randn(n, 64, 49, device="cuda").bfloat16()
y = torch.randn(64, 1, device="cuda").bfloat16()
pb.fused_elemwise_add_activation(x, y, ["relu", "elementwise_add"])   # bcast_ragged_rowscalar_kernel
pb.less_than(x, y)                                                    # map_ragged_rowscalar_kernel
torch.cuda.synchronize()
print("ok")
