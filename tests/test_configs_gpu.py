"""The five BASELINE.json configs at their full sizes: direct comparison with the
oracle where it finishes in seconds, plus size-independent properties (closed-form
column sums, linearity, idempotence, checksum of checksums)."""
import numpy as np
import pytest
import torch

import gpu_util as G
import paddle2_b200 as pb

pytestmark = pytest.mark.gpu


def test_c1_sum_fp32_32x1024x4096(oracle):
    g = torch.Generator(device="cuda").manual_seed(123)
    x = torch.rand(32, 1024, 4096, device="cuda", generator=g)
    y = pb.sum(x, axis=-1)
    assert tuple(y.shape) == (32, 1024) and y.dtype == torch.float32
    xh = x.cpu().numpy()
    want = oracle.reduce("sum", xh, [-1])
    np.testing.assert_allclose(G.to_np(y), want, rtol=G.FP32_RTOL)
    # checksum of checksums: sum of the row sums == full sum
    total = pb.sum(y)
    np.testing.assert_allclose(float(total), float(pb.sum(x)), rtol=G.FP32_RTOL)
    np.testing.assert_allclose(float(total), xh.astype(np.float64).sum(), rtol=G.FP32_RTOL)
    # linearity: sum(2x) == 2 sum(x) bit for bit (scaling by 2 is exact)
    assert torch.equal(pb.sum(pb.scale(x, 2.0), axis=-1), pb.scale(y, 2.0))
    # keepdim changes the shape only
    assert torch.equal(pb.sum(x, axis=-1, keepdim=True).squeeze(-1), y)


def test_c2_broadcast_add_relu_256x64x112x112(oracle):
    g = torch.Generator(device="cuda").manual_seed(123)
    x = torch.rand(256, 64, 112, 112, device="cuda", generator=g) * 2 - 1
    yb = torch.rand(64, 1, 1, device="cuda", generator=g) * 2 - 1
    out = pb.fused_elemwise_add_activation(x, yb, ["relu", "elementwise_add"])
    assert out.shape == x.shape
    # oracle on a slab (the CPU walk over all 205M elements is the slow part, not the GPU)
    xs = x[:4].cpu().numpy()
    want = oracle.elementwise("add", [xs, yb.cpu().numpy()], act="relu")
    assert np.array_equal(G.to_np(out[:4]), want)
    assert np.array_equal(G.to_np(out[-1:]), oracle.elementwise("add", [x[-1:].cpu().numpy(), yb.cpu().numpy()], act="relu"))
    # properties over the whole tensor: idempotence of relu, non-negativity, agreement with the unfused pair
    assert bool(pb.all(pb.maximum(out, pb.scale(out, 0.0)) == out))
    assert float(pb.min(out)) >= 0.0
    assert torch.equal(pb.relu(pb.add(x, yb)), out)
    assert torch.equal(pb.relu(out), out)
    # per-channel identity: where x + y > 0 the output is exactly x + y
    s = pb.add(x, yb)
    assert torch.equal(torch.where(s > 0, s, torch.zeros_like(s)), out)


def test_c3_argmax_max_int64_and_bf16_8192x32000(oracle):
    rows, n = 8192, 32000
    g = torch.Generator(device="cuda").manual_seed(123)
    x = torch.randint(-(2 ** 40), 2 ** 40, (rows, n), device="cuda", generator=g, dtype=torch.int64)
    # duplicate the row maximum at a later column in every 8th row (forces the tie-break)
    mx, am = torch.max(x, dim=1)
    dup_rows = torch.arange(0, rows, 8, device="cuda")
    dup_cols = torch.clamp(am[dup_rows] + 17, max=n - 1)
    x[dup_rows, dup_cols] = mx[dup_rows]
    idx, val = pb.argmax_with_value(x, axis=-1)
    xh = x.cpu().numpy()
    assert np.array_equal(G.to_np(idx), oracle.argreduce("argmax", xh, -1))
    assert np.array_equal(G.to_np(val), oracle.reduce("max", xh, [-1]))
    assert torch.equal(pb.argmax(x, axis=-1), idx) and torch.equal(pb.max(x, axis=-1), val)
    # property: x[row, idx] == max and nothing earlier equals it
    assert torch.equal(x.gather(1, idx[:, None])[:, 0], val)
    assert np.array_equal(G.to_np(pb.argmin(x, axis=-1)), oracle.argreduce("argmin", xh, -1))
    del x, xh
    xb = torch.randn(rows, n, device="cuda", generator=g).bfloat16()
    x32 = xb.float().cpu().numpy()
    idx, val = pb.argmax_with_value(xb, axis=-1)
    assert np.array_equal(G.to_np(idx), oracle.argreduce("argmax", x32, -1))
    assert np.array_equal(G.to_np(val), oracle.reduce("max", x32, [-1]))
    assert torch.equal(pb.max(xb, axis=-1), val)


def test_c4_row_mean_var_bf16_16384x8192(oracle):
    g = torch.Generator(device="cuda").manual_seed(123)
    xb = (torch.randn(16384, 8192, device="cuda", generator=g) * (2 ** 0.5) + 0.5).bfloat16()
    m, v = pb.row_mean_var(xb)
    x32 = xb.float().cpu().numpy()
    mean, var = oracle.rowstats_f32(x32, unbiased=True)
    np.testing.assert_allclose(G.to_np(m), mean, rtol=G.BF16_RTOL, atol=2e-3)
    np.testing.assert_allclose(G.to_np(v), var, rtol=G.BF16_RTOL)
    m32, v32 = pb.row_mean_var(xb, out_dtype="float32")
    np.testing.assert_allclose(G.to_np(m32), mean, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(G.to_np(v32), var, rtol=1e-4)
    # shift invariance of the variance: var(x) == var(x + c) up to fp32 rounding of the shift
    xs = (xb.float() + 0.0).bfloat16()
    assert torch.equal(pb.row_mean_var(xs)[1], v)
    # mean alone agrees with the reduce kernel's mean
    np.testing.assert_allclose(G.to_np(pb.mean(xb, axis=-1)), G.to_np(m), rtol=G.BF16_RTOL, atol=2e-3)


def _closed_form_shard(rows, cols, row0=0, chunk=8192):
    """x[r, c] = ((131 r + 7 c) mod 1024) / 1024: every run of 1024 rows sums to 511.5
    per column, so column sums are known exactly without a host copy."""
    x = torch.empty(rows, cols, device="cuda", dtype=torch.float32)
    c = torch.arange(cols, device="cuda", dtype=torch.int64)[None, :] * 7
    for r0 in range(0, rows, chunk):
        r = torch.arange(row0 + r0, row0 + min(r0 + chunk, rows), device="cuda", dtype=torch.int64)[:, None] * 131
        x[r0:r0 + r.shape[0]] = ((r + c) % 1024).float() / 1024.0
    return x


def test_c5_global_sum_axis0_full_shard(oracle):
    rows, cols = 524288, 4096          # 8 GiB of fp32 = 2^31 elements (numel > INT32_MAX)
    x = _closed_form_shard(rows, cols)
    assert x.numel() == 2 ** 31
    y = pb.sum(x, axis=0)
    want = np.full(cols, (rows // 1024) * 511.5, dtype=np.float64)
    np.testing.assert_allclose(G.to_np(y).astype(np.float64), want, rtol=G.FP32_RTOL)
    assert torch.equal(pb.sum(x, axis=0), y)                    # deterministic finalise
    # oracle on a 64 Ki-row slab of the same tensor.  The reference's CPU kernel adds the
    # 65536 rows of a column one after another in fp32 (GenericDimReducer), so ITS result is
    # off the exact sum by more than 1e-5 here; parity is stated as a triangle: the GPU is
    # within 1e-5 of the exact (fp64) sum, and no further from the reference than the
    # reference is from the exact sum plus that tolerance.
    slab = x[:65536].cpu().numpy()
    ref = oracle.reduce("sum", slab, [0]).astype(np.float64)
    exact = slab.astype(np.float64).sum(0)
    got = G.to_np(pb.sum(x[:65536], axis=0)).astype(np.float64)
    assert np.all(np.abs(got - exact) <= G.FP32_RTOL * exact)
    assert np.all(np.abs(got - ref) <= np.abs(ref - exact) + G.FP32_RTOL * exact)
    # full reduction of 2^31 elements through the split kernel; row reduce at 64-bit offsets
    np.testing.assert_allclose(float(pb.sum(x)), want.sum(), rtol=G.FP32_RTOL)
    rs = pb.sum(x, axis=1)
    np.testing.assert_allclose(float(pb.sum(rs)), want.sum(), rtol=G.FP32_RTOL)
    np.testing.assert_allclose(G.to_np(rs[-3:]), oracle.reduce("sum", x[-3:].cpu().numpy(), [1]), rtol=G.FP32_RTOL)
    np.testing.assert_allclose(float(pb.max(x)), 1023.0 / 1024.0)
    np.testing.assert_allclose(G.to_np(pb.mean(x, axis=0)).astype(np.float64), want / rows, rtol=G.FP32_RTOL)


def test_c5_sharded_reducer_single_rank(oracle):
    """world_size 1: b200phi_sharded_reduce with no communicator is the local reduce."""
    import os
    import torch.distributed as dist
    from paddle2_b200 import _lib as L
    from paddle2_b200.dist import ShardedReducer
    created = False
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        dist.init_process_group("gloo", rank=0, world_size=1)
        created = True
    try:
        red = ShardedReducer()
        assert red.comm is None
        x = torch.rand(20000, 512, device="cuda")
        assert torch.equal(red.sum(x), pb.sum(x, axis=0))
        np.testing.assert_allclose(G.to_np(red.reduce(L.MEAN, x)), G.to_np(pb.mean(x, axis=0)))
        red.close()
    finally:
        if created:
            dist.destroy_process_group()
