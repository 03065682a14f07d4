"""One launch of each kernel added in round 2 (plus the C2 / C3a kernels VERDICT asked digests for),
for an ncu capture:  ncu --set full -k regex:'tile_|arg_row_split|arg_col_split|bcast_rows|arg_row_kernel' ...
Set B200PHI_NO_TWO_STAGE=1 to keep the general patterns on the single-pass tiled kernels."""
import sys
import torch
sys.path.insert(0, '.')
import paddle2_b200 as pb

g = torch.Generator(device="cuda").manual_seed(1)
x = torch.rand(256, 64, 112, 112, device="cuda", generator=g)
pb.sum(x, axis=[0, 2])                       # tile_keep_kernel (or two-stage: col + col)
pb.sum(x, axis=[1, 3])                       # tile_red_kernel  (or two-stage: row + col)
y = torch.rand(64, 1, 1, device="cuda") * 2 - 1
pb.fused_elemwise_add_activation(x, y, ["relu", "elementwise_add"])   # C2: bcast_rows_kernel
xb = x.bfloat16()
pb.sum(xb, axis=[0, 2]); pb.sum(xb, axis=[1, 3])
del x, xb
x = torch.rand(32, 64, 48, 80, 6, 10, device="cuda", generator=g)
pb.min(x, axis=[2, 4])
del x
x = torch.randn(1 << 30, device="cuda", generator=g).bfloat16()
pb.argmax(x)                                 # arg_row_split_kernel
del x
x = torch.randn(131072, 4096, device="cuda", generator=g)
pb.argmax(x, axis=0)                         # arg_col_split_kernel
del x
x = torch.randint(-(2 ** 40), 2 ** 40, (8192, 32000), device="cuda", dtype=torch.int64)
pb.argmax_with_value(x, axis=-1)             # C3a: arg_row_kernel<long long>
torch.cuda.synchronize()
print("ok")
