"""The issue-bound shapes VERDICT r1 #7 names: 16-bit rows that are not whole 16-byte vectors (sum /
max / logsumexp / argmax / mean+var on 101..1001 columns, statistics on unaligned 16-24 KiB rows) and
ragged broadcasts (rows of 49 / 196 elements against a per-channel operand).  ~2-4 GiB tensors,
CUDA events over 3 calls through the public ops."""
import sys, torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import ops


def t(fn, n=3):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


N = 1 << 30
SHORT = len(sys.argv) > 1 and sys.argv[1] == "short"   # only the short-row section (knob sweeps)
STATS = len(sys.argv) > 1 and sys.argv[1] == "stats"   # only the long statistics rows
for cols in (() if STATS else (101, 201, 301, 501, 1001)):
    rows = N // cols
    x = torch.randn(rows, cols, device="cuda").bfloat16()
    for name, fn in (("sum", lambda: pb.sum(x, axis=-1)), ("max", lambda: pb.max(x, axis=-1)),
                     ("logsumexp", lambda: ops.logsumexp(x, -1)), ("argmax", lambda: pb.argmax(x, axis=-1)),
                     ("mean+var", lambda: pb.row_mean_var(x))):
        us = t(fn)
        print(f"{name:10s} bf16 [{rows},{cols}] {us:9.1f} us {x.numel()*2/us/1e6:6.2f} TB/s", flush=True)
    del x
if SHORT:
    sys.exit(0)
for cols in (8193, 9001, 10001, 12289):
    rows = N // cols
    x = torch.randn(rows, cols, device="cuda").bfloat16()
    us = t(lambda: pb.row_mean_var(x))
    print(f"mean+var   bf16 [{rows},{cols}] {cols*2:6d} B/row {us:9.1f} us {x.numel()*2/us/1e6:6.2f} TB/s", flush=True)
    del x
for cols in (4097, 5001, 6001, 7001, 8191):
    rows = N // cols
    x = torch.randn(rows, cols, device="cuda")
    us = t(lambda: pb.row_mean_var(x))
    print(f"mean+var   fp32 [{rows},{cols}] {cols*4:6d} B/row {us:9.1f} us {x.numel()*4/us/1e6:6.2f} TB/s", flush=True)
    del x
if STATS:
    sys.exit(0)
for dt in (torch.bfloat16, torch.float32):
    for hw in (49, 196):
        n = (1 << 30) // (64 * hw)
        x = torch.randn(n, 64, hw, device="cuda").to(dt)
        y = torch.randn(64, 1, device="cuda").to(dt)
        us = t(lambda: pb.fused_elemwise_add_activation(x, y, ["relu", "elementwise_add"]))
        print(f"relu(x+y)  {str(dt)[6:]:8s} [{n},64,{hw}]+[64,1] {us:9.1f} us {2*x.numel()*x.element_size()/us/1e6:6.2f} TB/s", flush=True)
        us = t(lambda: pb.less_than(x, y))
        print(f"less_than  {str(dt)[6:]:8s} [{n},64,{hw}] vs [64,1] {us:9.1f} us {(x.numel()*x.element_size()+x.numel())/us/1e6:6.2f} TB/s", flush=True)
        del x
