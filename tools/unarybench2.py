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
for dt in (torch.float32, torch.bfloat16):
    es = 4 if dt == torch.float32 else 2
    x = (torch.rand((1 << 30) // es, device="cuda") + 0.5).to(dt).view(-1, 4096)
    y = torch.randn(4096, device="cuda").to(dt)
    line = f"{str(dt)[6:]:9s}"
    fams = [("log", lambda: ops.log(x)), ("sqrt", lambda: ops.sqrt(x)), ("rsqrt", lambda: ops.rsqrt(x)), ("recip", lambda: ops.reciprocal(x)),
            ("swish", lambda: ops.swish(x)), ("mish", lambda: ops.mish(x)), ("softplus", lambda: ops.softplus(x)), ("elu", lambda: ops.elu(x)),
            ("logsig", lambda: ops.log_sigmoid(x)), ("softsign", lambda: ops.softsign(x)), ("expm1", lambda: ops.expm1(x)),
            ("divide", lambda: pb.divide(x, y)),
            ("gelu(x+y)", lambda: pb.fused_elemwise_add_activation(x, y, ["gelu", "elementwise_add"])),
            ("sigm(x+y)", lambda: pb.fused_elemwise_add_activation(x, y, ["sigmoid", "elementwise_add"])),
            ("tanh(x+y)", lambda: pb.fused_elemwise_add_activation(x, y, ["tanh", "elementwise_add"]))]
    for name, fn in fams:
        try:
            us = t(fn)
            line += f"  {name} {2 * x.numel() * es / us / 1e6:4.2f}"
        except Exception as e:
            line += f"  {name} ERR"
    print(line, flush=True)
