"""Short rows that are not whole 16-byte vectors (classifier widths, 7x7 maps, odd hidden sizes),
every family that reduces along the row, 1-2 GiB tensors, CUDA events over 3 calls.  Run once as is
and once with B200PHI_NO_ROW_STAGED=1 to compare the thread-per-row shared-memory kernels
(rows_staged.cuh) with the sub-warp kernels they replace.  argv[1] = comma list of column counts."""
import sys
import torch
sys.path.insert(0, '.')
import paddle2_b200 as pb
from paddle2_b200 import ops


def t(fn, n=3):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


COLS = [int(c) for c in sys.argv[1].split(",")] if len(sys.argv) > 1 else [3, 7, 10, 21, 33, 49, 101, 127]
ALIGNED_TOO = len(sys.argv) > 2 and sys.argv[2] == "all"   # also rows that are whole 16-byte vectors
NBYTES = 1 << 30
for dt in (torch.float32, torch.bfloat16, torch.int64):
    es = torch.empty((), dtype=dt).element_size()
    for cols in COLS:
        if (cols * es % 16 == 0 and not ALIGNED_TOO) or cols * es > 1024:
            continue
        rows = NBYTES // (cols * es)
        if dt == torch.int64:
            x = torch.randint(-1000, 1000, (rows, cols), device="cuda", dtype=dt)
            fams = (("sum", lambda: pb.sum(x, axis=-1)), ("max", lambda: pb.max(x, axis=-1)),
                    ("argmax", lambda: pb.argmax(x, axis=-1)))
        else:
            x = torch.randn(rows, cols, device="cuda").to(dt)
            fams = (("sum", lambda: pb.sum(x, axis=-1)), ("max", lambda: pb.max(x, axis=-1)),
                    ("argmax", lambda: pb.argmax(x, axis=-1)), ("mean+var", lambda: pb.row_mean_var(x)),
                    ("logsumexp", lambda: ops.logsumexp(x, -1)))
        line = f"{str(dt)[6:]:9s} [{rows},{cols}] {cols * es:4d} B/row "
        for name, fn in fams:
            us = t(fn)
            line += f" {name} {x.numel() * es / us / 1e6:5.2f}"
        print(line + "  TB/s", flush=True)
        del x
