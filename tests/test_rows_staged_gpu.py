"""Short rows that are not whole aligned 16-byte vectors (classifier widths, odd hidden sizes):
rows_staged_kernel -- spans of 128 rows bulk-copied into shared memory, a thread per row -- for the
reduce family, argmax / argmin, row statistics and logsumexp, against the oracle.  The shapes
exercise a last span that is not full, a tensor whose first / last bytes are not 16-byte aligned
(the byte-copied edges), every element size, and ties for the arg kernels."""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_util as G
import paddle2_b200 as pb
from paddle2_b200 import _lib as L
from paddle2_b200 import ops

pytestmark = pytest.mark.gpu

STAGED = 14


def staged_expected(x):
    """The plain reductions leave rows under 16 / 32 / 64 bytes (2- / 4- / 8-byte elements) to the
    sub-warp kernels."""
    row_bytes = x.shape[-1] * x.element_size()
    es = x.element_size()
    return row_bytes >= (16 if es <= 2 else (32 if es == 4 else 64))


def plan_kind(x, op=L.SUM):
    d = L.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = op, ops._phi_dtype(x.dtype), 0, x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.naxes = 1
    d.axes[0] = x.dim() - 1
    d.x = x.data_ptr()
    return L.lib().b200phi_reduce_plan_kind(C.byref(d))


def offset_view(flat, rows, cols, off):
    """[rows, cols] view that starts `off` elements into a fresh allocation (an unaligned base)."""
    buf = torch.empty(rows * cols + off + 16, dtype=flat.dtype, device="cuda")
    v = buf[off:off + rows * cols].view(rows, cols)
    v.copy_(flat.view(rows, cols))
    return v


# (rows, cols): row bytes 2..384, never a multiple of 16
SHAPES_16 = [(5000, 2), (4096, 3), (7777, 33), (20001, 101), (4100, 191)]
SHAPES_32 = [(6000, 2), (4096, 7), (9999, 10), (5003, 33), (4500, 95)]
SHAPES_64 = [(5000, 3), (4097, 47)]


@pytest.mark.parametrize("rows,cols", SHAPES_32)
@pytest.mark.parametrize("off", [0, 1, 3])
@pytest.mark.parametrize("op", ["sum", "mean", "max", "min", "prod"])
def test_fp32_reduce(oracle, rows, cols, off, op):
    rng = np.random.default_rng(rows + cols + off)
    x = rng.random((rows, cols)).astype(np.float32)
    if op == "prod":
        x = (0.9 + 0.2 * x).astype(np.float32)
    xt = offset_view(G.to_dev(x).flatten(), rows, cols, off)
    assert (plan_kind(xt, getattr(L, op.upper())) == STAGED) == staged_expected(xt)
    got = G.to_np(getattr(pb, op)(xt, axis=-1))
    want = oracle.reduce(op, x, [-1])
    G.assert_reduce_close(got, want, x, [-1], op, max(G.FP32_RTOL, cols * 2.0 ** -24 if op == "prod" else 0))


@pytest.mark.parametrize("rows,cols", SHAPES_16)
@pytest.mark.parametrize("off", [0, 1, 5])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_16bit_reduce_arg_stats_lse(oracle, rows, cols, off, dt):
    rng = np.random.default_rng(rows * 3 + cols + off)
    x32 = torch.from_numpy(rng.standard_normal((rows, cols)).astype(np.float32)).to(dt).float().numpy()
    xt = offset_view(torch.from_numpy(x32).cuda().to(dt).flatten(), rows, cols, off)
    assert (plan_kind(xt) == STAGED) == staged_expected(xt)
    tol = G.BF16_RTOL
    np.testing.assert_allclose(pb.sum(xt, axis=-1).float().cpu().numpy(), oracle.reduce("sum", x32, [-1]),
                               rtol=tol, atol=tol * np.abs(x32).sum(-1).max())
    np.testing.assert_allclose(pb.sum(xt, axis=-1, dtype="float32").cpu().numpy(), oracle.reduce("sum", x32, [-1]),
                               rtol=1e-5, atol=1e-5 * np.abs(x32).sum(-1).max())
    assert np.array_equal(pb.max(xt, axis=-1).float().cpu().numpy(), oracle.reduce("max", x32, [-1]))
    assert np.array_equal(pb.min(xt, axis=-1).float().cpu().numpy(), oracle.reduce("min", x32, [-1]))
    # 16-bit values tie often: lowest index
    idx, val = pb.argmax_with_value(xt, axis=-1)
    assert np.array_equal(idx.cpu().numpy(), oracle.argreduce("argmax", x32, -1))
    assert np.array_equal(val.float().cpu().numpy(), oracle.reduce("max", x32, [-1]))
    assert np.array_equal(pb.argmin(xt, axis=-1, dtype="int32").cpu().numpy(),
                          oracle.argreduce("argmin", x32, -1).astype(np.int32))
    if cols > 1:
        m, v = pb.row_mean_var(xt, out_dtype="float32")
        mean, var = oracle.rowstats_f32(x32, unbiased=True)
        np.testing.assert_allclose(m.cpu().numpy(), mean, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(v.cpu().numpy(), var, rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(ops.logsumexp(xt, -1).float().cpu().numpy(), oracle.logsumexp(x32, [-1]),
                               rtol=tol)


@pytest.mark.parametrize("rows,cols", SHAPES_64 + SHAPES_32[:3])
@pytest.mark.parametrize("off", [0, 1])
def test_integers_bit_exact(oracle, rows, cols, off):
    rng = np.random.default_rng(rows + 7 * cols + off)
    for npdt in (np.int64, np.int32, np.int16, np.uint8, np.int8):
        if np.dtype(npdt).itemsize * cols % 16 == 0 or np.dtype(npdt).itemsize * cols < 2:
            continue
        x = rng.integers(0 if npdt == np.uint8 else -100, 100, size=(rows, cols)).astype(npdt)
        xt = offset_view(G.to_dev(x).flatten(), rows, cols, off)
        assert (plan_kind(xt) == STAGED) == staged_expected(xt)
        # int32 sums promote to int64 (SumRawInferMeta); the other integers keep their type and wrap
        sum_dt = np.int64 if npdt in (np.int32, np.int64) else npdt
        assert np.array_equal(G.to_np(pb.sum(xt, axis=-1)), oracle.reduce("sum", x, [-1], out_dtype=sum_dt))
        assert np.array_equal(G.to_np(pb.max(xt, axis=-1)), oracle.reduce("max", x, [-1]))
        if npdt in (np.int64, np.int32, np.int16, np.uint8):
            assert np.array_equal(G.to_np(pb.argmax(xt, axis=-1)), oracle.argreduce("argmax", x, -1))
            assert np.array_equal(G.to_np(pb.argmin(xt, axis=-1)), oracle.argreduce("argmin", x, -1))
    xb = rng.random((rows, cols)) > 0.9
    xt = offset_view(G.to_dev(xb).flatten(), rows, cols, off)
    assert np.array_equal(G.to_np(pb.any(xt, axis=-1)), oracle.reduce("any", xb, [-1]))
    assert np.array_equal(G.to_np(pb.all(offset_view(G.to_dev(~xb).flatten(), rows, cols, off), axis=-1)),
                          oracle.reduce("all", ~xb, [-1]))


def test_fp64_and_fp32_stats_lse(oracle):
    rng = np.random.default_rng(5)
    xd = rng.standard_normal((4321, 13))
    xt = offset_view(G.to_dev(xd).flatten(), 4321, 13, 1)
    np.testing.assert_allclose(G.to_np(pb.sum(xt, axis=-1)), oracle.reduce("sum", xd, [-1]), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(G.to_np(ops.logsumexp(xt, -1)), oracle.logsumexp(xd, [-1]), rtol=1e-11)
    xf = rng.standard_normal((8191, 21)).astype(np.float32)
    xt = offset_view(G.to_dev(xf).flatten(), 8191, 21, 2)
    m, v = pb.row_mean_var(xt)
    mean, var = oracle.rowstats_f32(xf, unbiased=True)
    np.testing.assert_allclose(m.cpu().numpy(), mean, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(v.cpu().numpy(), var, rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(G.to_np(ops.logsumexp(xt, -1)), oracle.logsumexp(xf, [-1]), rtol=1e-5, atol=1e-6)
    # all -inf row and a row with one huge element
    xf[5, :] = -np.inf
    xf[6, 3] = 80.0
    xt = offset_view(G.to_dev(xf).flatten(), 8191, 21, 2)
    got = G.to_np(ops.logsumexp(xt, -1))
    assert got[5] == -np.inf
    np.testing.assert_allclose(got[6], oracle.logsumexp(xf, [-1])[6], rtol=1e-5)


def test_large_and_canaries():
    """Many spans per CTA (the ring wraps many times), outputs framed by sentinels."""
    rows, cols = 3_000_001, 37
    x = torch.randn(rows * cols + 3, device="cuda")[3:].view(rows, cols)
    frame = torch.full((rows + 64,), 12345.0, device="cuda")
    out = frame[32:32 + rows]
    pb.sum(x, axis=-1, out=out)
    want = x.double().sum(-1)
    assert torch.all((out.double() - want).abs() <= 1e-5 * x.double().abs().sum(-1) + 1e-30)
    assert torch.all(frame[:32] == 12345.0) and torch.all(frame[32 + rows:] == 12345.0)
    a = pb.sum(x, axis=-1)
    assert torch.equal(a, pb.sum(x, axis=-1))
    idx = pb.argmax(x, axis=-1)
    assert torch.equal(idx, x.argmax(-1))


# ---- rows that ARE whole 16-byte vectors: visited from staggered starts, wrapped around ---------
ALIGNED_16 = [(5000, 8), (4096, 16), (6001, 24), (4500, 64), (4099, 96), (5000, 128), (4097, 256)]
ALIGNED_32 = [(5000, 4), (4096, 8), (6001, 24), (4500, 40), (4099, 64), (4200, 128), (4097, 208)]


@pytest.mark.parametrize("rows,cols", ALIGNED_16)
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_aligned_16bit_rows(oracle, rows, cols, dt):
    rng = np.random.default_rng(rows + cols)
    # few distinct values: every row has its extreme several times, on both sides of the wrap
    x32 = rng.integers(-3, 4, size=(rows, cols)).astype(np.float32)
    x32[7, :] = -np.inf           # nothing beats the identity: index 0
    x32[8, :] = 2.0               # all equal: index 0
    xt = torch.from_numpy(x32).cuda().to(dt)
    idx, val = pb.argmax_with_value(xt, axis=-1)
    assert np.array_equal(idx.cpu().numpy(), oracle.argreduce("argmax", x32, -1))
    assert np.array_equal(val.float().cpu().numpy()[9:], oracle.reduce("max", x32, [-1])[9:])
    assert np.array_equal(pb.argmin(xt, axis=-1).cpu().numpy()[9:], oracle.argreduce("argmin", x32, -1)[9:])
    xr = torch.from_numpy(rng.standard_normal((rows, cols)).astype(np.float32)).to(dt)
    xr32 = xr.float().numpy()
    xd = xr.cuda()
    m, v = pb.row_mean_var(xd, out_dtype="float32")
    mean, var = oracle.rowstats_f32(xr32, unbiased=True)
    np.testing.assert_allclose(m.cpu().numpy(), mean, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(v.cpu().numpy(), var, rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(ops.logsumexp(xd, -1).float().cpu().numpy(), oracle.logsumexp(xr32, [-1]),
                               rtol=G.BF16_RTOL)
    np.testing.assert_allclose(pb.sum(xd, axis=-1, dtype="float32").cpu().numpy(), oracle.reduce("sum", xr32, [-1]),
                               rtol=1e-5, atol=1e-5 * np.abs(xr32).sum(-1).max())
    assert np.array_equal(pb.max(xd, axis=-1).float().cpu().numpy(), oracle.reduce("max", xr32, [-1]))
    assert np.array_equal(idx.cpu().numpy()[:7], oracle.argreduce("argmax", x32, -1)[:7])


@pytest.mark.parametrize("rows,cols", ALIGNED_32)
def test_aligned_fp32_int_rows(oracle, rows, cols):
    rng = np.random.default_rng(rows * 5 + cols)
    xi = rng.integers(-2, 3, size=(rows, cols)).astype(np.int32)
    xt = G.to_dev(xi)
    assert np.array_equal(G.to_np(pb.argmax(xt, axis=-1)), oracle.argreduce("argmax", xi, -1))
    assert np.array_equal(G.to_np(pb.argmin(xt, axis=-1, dtype="int32")), oracle.argreduce("argmin", xi, -1).astype(np.int32))
    assert np.array_equal(G.to_np(pb.sum(xt, axis=-1)), oracle.reduce("sum", xi, [-1], out_dtype=np.int64))
    if cols % 2 == 0:
        xl = rng.integers(-2, 3, size=(rows, cols // 2)).astype(np.int64)
        assert np.array_equal(G.to_np(pb.argmax(G.to_dev(xl), axis=-1)), oracle.argreduce("argmax", xl, -1))
        assert np.array_equal(G.to_np(pb.min(G.to_dev(xl), axis=-1)), oracle.reduce("min", xl, [-1]))
    xf = rng.standard_normal((rows, cols)).astype(np.float32)
    xf[3, :] = -np.inf
    xd = G.to_dev(xf)
    assert np.array_equal(G.to_np(pb.argmax(xd, axis=-1)), oracle.argreduce("argmax", xf, -1))
    got = G.to_np(ops.logsumexp(xd, -1))
    want = oracle.logsumexp(xf, [-1])
    assert got[3] == -np.inf
    np.testing.assert_allclose(np.delete(got, 3), np.delete(want, 3), rtol=1e-5, atol=1e-6)
    xf[3, :] = 1.0
    xd = G.to_dev(xf)
    m, v = pb.row_mean_var(xd)
    mean, var = oracle.rowstats_f32(xf, unbiased=True)
    np.testing.assert_allclose(m.cpu().numpy(), mean, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(v.cpu().numpy(), var, rtol=1e-5, atol=1e-7)
    for op in ("sum", "max", "min", "mean"):
        got = G.to_np(getattr(pb, op)(xd, axis=-1))
        G.assert_reduce_close(got, oracle.reduce(op, xf, [-1]), xf, [-1], op, G.FP32_RTOL)
