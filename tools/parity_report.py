"""Measured distance between the GPU results, the reference's CPU arithmetic and the exact (fp64)
value for the five BASELINE.json configs (state the numbers, not only the bound).

For every config: max |gpu - ref| / |ref| and max |gpu - exact| / |exact| over the outputs (plus the
reference's own |ref - exact| / |exact|), `ref` = oracle (the reference's own functors where
oracle/_ref exists, else the bit-identical C port) on the full tensor where that takes seconds and on
a stated slab otherwise.  Integer / index results are compared bit for bit.  Prints one JSON object."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import paddle2_b200 as pb
from oracle import pyoracle as O

IMPL = "ref" if O.have_ref() else "port"


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))


def main():
    out = {"oracle": "reference functors (oracle/_ref)" if IMPL == "ref" else "C port (bit-identical to the reference build)"}
    g = torch.Generator(device="cuda").manual_seed(123)
    # C1
    x = torch.rand(32, 1024, 4096, device="cuda", generator=g)
    y = pb.sum(x, axis=-1).cpu().numpy()
    xh = x.cpu().numpy()
    ref = O.reduce("sum", xh, [-1], impl=IMPL)
    exact = xh.astype(np.float64).sum(-1)
    out["C1 sum fp32 [32,1024,4096] axis -1"] = {
        "gpu_vs_ref": rel(y, ref), "gpu_vs_exact": rel(y, exact), "ref_vs_exact": rel(ref, exact),
        "tolerance": 1e-5, "compared_on": "the full tensor"}
    del x, xh
    # C2 (one IEEE add + a select: bit-exact)
    x = torch.rand(256, 64, 112, 112, device="cuda", generator=g) * 2 - 1
    yb = torch.rand(64, 1, 1, device="cuda", generator=g) * 2 - 1
    o = pb.fused_elemwise_add_activation(x, yb, ["relu", "elementwise_add"])
    xs = x[:16].cpu().numpy()
    want = O.elementwise("add", [xs, yb.cpu().numpy()], act="relu", impl=IMPL)
    got = o[:16].cpu().numpy()
    out["C2 relu(x+y) fp32 [256,64,112,112]+[64,1,1]"] = {
        "bit_exact": bool(np.array_equal(got, want)), "gpu_vs_ref": rel(got + 1.0, want + 1.0),
        "compared_on": "a [16,64,112,112] slab against the oracle; the whole tensor against relu(add(x, y)) on the GPU",
        "whole_tensor_equals_unfused": bool(torch.equal(pb.relu(pb.add(x, yb)), o))}
    del x, o
    # C3
    x = torch.randint(-(2 ** 40), 2 ** 40, (8192, 32000), device="cuda", generator=g, dtype=torch.int64)
    idx, val = pb.argmax_with_value(x, axis=-1)
    xh = x.cpu().numpy()
    out["C3a argmax+max int64 [8192,32000]"] = {
        "indices_bit_exact": bool(np.array_equal(idx.cpu().numpy(), O.argreduce("argmax", xh, -1, impl=IMPL))),
        "values_bit_exact": bool(np.array_equal(val.cpu().numpy(), O.reduce("max", xh, [-1], impl=IMPL))),
        "compared_on": "the full tensor"}
    del x, xh
    xb = torch.randn(8192, 32000, device="cuda", generator=g).bfloat16()
    x32 = xb.float().cpu().numpy()
    idx, val = pb.argmax_with_value(xb, axis=-1)
    out["C3b argmax+max bf16 [8192,32000]"] = {
        "indices_bit_exact": bool(np.array_equal(idx.cpu().numpy(), O.argreduce("argmax", x32, -1, impl=IMPL))),
        "values_bit_exact": bool(np.array_equal(val.float().cpu().numpy(), O.reduce("max", x32, [-1], impl=IMPL))),
        "compared_on": "the full tensor widened to fp32 (phi registers no bf16 CPU argmax / max)"}
    del xb, x32
    # C4
    xb = (torch.randn(16384, 8192, device="cuda", generator=g) * (2 ** 0.5) + 0.5).bfloat16()
    x32 = xb.float().cpu().numpy()
    m, v = pb.row_mean_var(xb)
    m32, v32 = pb.row_mean_var(xb, out_dtype="float32")
    mean_ref, var_ref = O.rowstats_f32(x32, unbiased=True)
    x64 = x32.astype(np.float64)
    out["C4 mean/var bf16 [16384,8192]"] = {
        "var_bf16_out_vs_ref": rel(v.float().cpu().numpy(), var_ref), "var_fp32_out_vs_ref": rel(v32.cpu().numpy(), var_ref),
        "var_fp32_out_vs_exact": rel(v32.cpu().numpy(), x64.var(axis=1, ddof=1)),
        "mean_fp32_out_vs_exact_abs": float(np.max(np.abs(m32.cpu().numpy() - x64.mean(1)))),
        "tolerance": 1e-2, "compared_on": "the full tensor widened to fp32 (oracle: the two-pass composite in fp32)"}
    del xb, x32, x64
    # C5: closed form over the full 8 GiB shard; the reference on a 64 Ki-row slab
    rows, cols = 524288, 4096
    x = torch.empty(rows, cols, device="cuda")
    c = torch.arange(cols, device="cuda", dtype=torch.int64)[None, :] * 7
    for r0 in range(0, rows, 16384):
        r = torch.arange(r0, r0 + 16384, device="cuda", dtype=torch.int64)[:, None] * 131
        x[r0:r0 + 16384] = ((r + c) % 1024).float() / 1024.0
    y = pb.sum(x, axis=0).cpu().numpy()
    exact_full = np.full(cols, (rows // 1024) * 511.5)
    slab = x[:65536].cpu().numpy()
    ref = O.reduce("sum", slab, [0], impl=IMPL)
    exact = slab.astype(np.float64).sum(0)
    got = pb.sum(x[:65536], axis=0).cpu().numpy()
    out["C5 sum fp32 [524288,4096] axis 0"] = {
        "full_shard_gpu_vs_exact": rel(y, exact_full),
        "slab65536_gpu_vs_ref": rel(got, ref), "slab65536_gpu_vs_exact": rel(got, exact),
        "slab65536_ref_vs_exact": rel(ref, exact), "tolerance": 1e-5,
        "compared_on": "closed-form column sums over the full 8 GiB shard; the reference (which adds a column's "
                       "rows one after another in fp32) on rows [0, 65536)"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
