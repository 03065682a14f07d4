"""Every launch class of the short-row kernels (sub-warp groups of 8 / 16 lanes, multi-row warps,
row batches, flat and ragged broadcast walks) on row counts that do not fill the last CTA.
Integer results bit-exact; floating point against float64 numpy."""
import numpy as np
import pytest
import torch

import gpu_util as G
import paddle2_b200 as pb
from paddle2_b200 import ops

pytestmark = pytest.mark.gpu

# columns chosen so that the row is 17, 33, 65, 129, 257, 513 16-byte vectors: one per class
FP32_COLS = [132, 260, 516, 1028, 2052]       # 33, 65, 129, 257, 513 vectors
BF16_COLS = [136, 264, 520, 1032, 2056, 4104]  # 17, 33, 65, 129, 257, 513 vectors


@pytest.mark.parametrize("rows", [1, 63, 64, 777])
@pytest.mark.parametrize("cols", FP32_COLS)
def test_row_reduce_classes_fp32_and_int(oracle, rows, cols):
    rng = np.random.default_rng(rows * 7 + cols)
    x = rng.standard_normal((rows, cols)).astype(np.float32)
    dx = G.to_dev(x)
    np.testing.assert_allclose(G.to_np(pb.sum(dx, axis=-1)), x.astype(np.float64).sum(-1), rtol=1e-5, atol=1e-4)
    assert np.array_equal(G.to_np(pb.max(dx, axis=-1)), x.max(-1))
    assert np.array_equal(G.to_np(pb.min(dx, axis=-1)), x.min(-1))
    xi = rng.integers(-1000, 1000, size=(rows, cols)).astype(np.int32)
    assert np.array_equal(G.to_np(pb.sum(G.to_dev(xi), axis=-1)), xi.sum(-1, dtype=np.int64))
    assert np.array_equal(G.to_np(pb.argmax(G.to_dev(xi), axis=-1)), oracle.argreduce("argmax", xi, -1))


@pytest.mark.parametrize("rows", [1, 65, 1000])
@pytest.mark.parametrize("cols", BF16_COLS)
def test_row_ops_classes_bf16(rows, cols):
    gen = torch.Generator().manual_seed(rows + cols)
    x = (torch.randn((rows, cols), generator=gen) * 2 + 0.5).bfloat16()
    x64 = x.double()
    xd = x.cuda()
    assert torch.equal(pb.max(xd, axis=-1).cpu(), x.amax(-1))
    np.testing.assert_allclose(pb.sum(xd, axis=-1, dtype="float32").cpu().numpy(), x64.sum(-1).numpy(), rtol=1e-5, atol=1e-3)
    m, v = pb.row_mean_var(xd, out_dtype="float32")
    np.testing.assert_allclose(m.cpu().numpy(), x64.mean(-1).numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(v.cpu().numpy(), x64.var(-1, unbiased=True).numpy() if cols > 1 else 0, rtol=1e-4)
    lse = ops.logsumexp(xd, -1)
    np.testing.assert_allclose(lse.double().cpu().numpy(), torch.logsumexp(x64, -1).numpy(), rtol=1e-2)
    # argmax: lowest index among the ties bf16 rounding creates
    want = np.argmax(x.float().numpy(), axis=-1)
    assert np.array_equal(pb.argmax(xd, axis=-1).cpu().numpy(), want)
    assert np.array_equal(pb.argmin(xd, axis=-1).cpu().numpy(), np.argmin(x.float().numpy(), axis=-1))


@pytest.mark.parametrize("xs,ys", [((7, 5, 49), (5, 1)), ((3, 64, 196), (64, 1)), ((33, 7, 50), (7, 50)),
                                   ((129, 196), (196,)), ((65, 64), (64,)), ((300, 3, 1028), (3, 1)),
                                   ((17, 2052), (17, 1)), ((5, 3, 7, 9), (3, 1, 9)), ((1000, 6), (6,))])
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16, torch.int64])
def test_broadcast_walk_classes(xs, ys, dt):
    """whole-vector rows, ragged rows with a per-row scalar / a shared row / a general operand,
    tiny rows: add, fused relu, a compare and a ternary backward functor; all exact"""
    gen = torch.Generator().manual_seed(len(xs) * 100 + xs[-1])
    x = (torch.randn(xs, generator=gen) * 4).to(dt)
    y = (torch.randn(ys, generator=gen) * 4).to(dt)
    xd, yd = x.cuda(), y.cuda()
    assert torch.equal(pb.add(xd, yd).cpu(), x + y)
    assert torch.equal(pb.add(yd, xd).cpu(), y + x)
    assert torch.equal(pb.less_than(xd, yd).cpu(), x < y)
    assert torch.equal(pb.maximum(xd, yd).cpu(), torch.maximum(x, y))
    if dt != torch.int64:
        got = pb.fused_elemwise_add_activation(xd, yd, ["relu", "elementwise_add"]).cpu()
        assert torch.equal(got, torch.relu(x + y))
    g = (torch.randn(xs, generator=gen) * 2).to(dt)
    got = ops._map("max_grad_x", [xd, yd, g.cuda()]).cpu()
    assert torch.equal(got, g * (x > y).to(dt))


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,cols", [(64, 101), (1000, 7), (777, 129), (200, 257), (129, 513), (65, 1001), (70, 1023),
                                       (67, 4097), (3, 50257)])
def test_unaligned_rows(rows, cols, dt):
    """rows that are not whole 16-byte vectors: the scalar sub-warp kernels (<= 1024 columns), the
    head / vectors / tail peel above that"""
    gen = torch.Generator().manual_seed(rows * 3 + cols)
    x = (torch.randn((rows, cols), generator=gen) * 3).to(dt)
    x64, xd = x.double(), x.cuda()
    m, v = pb.row_mean_var(xd, out_dtype="float32")
    np.testing.assert_allclose(m.cpu().numpy(), x64.mean(-1).numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(v.cpu().numpy(), x64.var(-1, unbiased=True).numpy(), rtol=1e-4)
    lse = ops.logsumexp(xd, -1)
    np.testing.assert_allclose(lse.double().cpu().numpy(), torch.logsumexp(x64, -1).numpy(),
                               rtol=1e-5 if dt == torch.float32 else 1e-2)
    assert np.array_equal(pb.argmax(xd, axis=-1).cpu().numpy(), np.argmax(x.float().numpy(), axis=-1))
    assert np.array_equal(pb.argmin(xd, axis=-1).cpu().numpy(), np.argmin(x.float().numpy(), axis=-1))
    idx, val = pb.argmax_with_value(xd, axis=-1)
    assert torch.equal(val.cpu(), x.amax(-1))
    assert torch.equal(pb.max(xd, axis=-1).cpu(), x.amax(-1))
    assert torch.equal(pb.min(xd, axis=-1).cpu(), x.amin(-1))
    s = pb.sum(xd, axis=-1).double().cpu().numpy()
    tol = 1e-5 if dt == torch.float32 else 1e-2
    np.testing.assert_allclose(s, x64.sum(-1).numpy(), rtol=tol, atol=tol * float(x64.abs().sum(-1).max()))


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16, torch.int32, torch.uint8, torch.float64])
@pytest.mark.parametrize("rows,cols", [(64, 129), (333, 255), (200, 257), (129, 511), (65, 513), (4099, 1001), (70, 1023)])
def test_unaligned_short_row_reduce(rows, cols, dt):
    """the scalar row kernel of the reduce family (129..1023 columns, not whole vectors), also from a
    base address that is one element past a 16-byte boundary"""
    gen = torch.Generator().manual_seed(rows + cols)
    if dt.is_floating_point:
        flat = (torch.randn(rows * cols + 1, generator=gen) * 3).to(dt)
    else:
        flat = torch.randint(0, 100, (rows * cols + 1,), generator=gen).to(dt)
    fd = flat.cuda()
    for off in (0, 1):
        x = flat[off:off + rows * cols].view(rows, cols)
        xd = fd[off:off + rows * cols].view(rows, cols)
        assert torch.equal(pb.max(xd, axis=-1).cpu(), x.amax(-1))
        assert torch.equal(pb.min(xd, axis=-1).cpu(), x.amin(-1))
        got = pb.sum(xd, axis=-1).cpu()
        if dt.is_floating_point:
            tol = {torch.float32: 1e-5, torch.bfloat16: 1e-2, torch.float64: 1e-12}[dt]
            ref = x.double().sum(-1)
            np.testing.assert_allclose(got.double().numpy(), ref.numpy(), rtol=tol,
                                       atol=tol * float(x.double().abs().sum(-1).max()))
        else:
            # SumRawInferMeta: uint8 sums stay uint8 (wrap), int32 sums widen to int64
            assert torch.equal(got.long(), x.long().sum(-1) % 256 if dt == torch.uint8 else x.long().sum(-1))


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16, torch.int32, torch.uint8, torch.float64, torch.int64])
@pytest.mark.parametrize("rows,cols", [(64, 16), (1000, 32), (4099, 48), (777, 64), (513, 80), (300, 96), (65, 128),
                                       (100000, 16), (33333, 112)])
def test_tiny_aligned_row_reduce(rows, cols, dt):
    """aligned rows of 1..16 vectors (row_tiny_kernel, G = 1..16 lanes per row, 8 rows per group) and
    the 17..32-vector rows that now take the batch kernel; row counts that end inside a group"""
    gen = torch.Generator().manual_seed(rows * 7 + cols)
    if dt.is_floating_point:
        x = (torch.randn(rows, cols, generator=gen) * 3).to(dt)
    else:
        x = torch.randint(0, 100, (rows, cols), generator=gen).to(dt)
    xd = x.cuda()
    assert torch.equal(pb.max(xd, axis=-1).cpu(), x.amax(-1))
    assert torch.equal(pb.min(xd, axis=-1).cpu(), x.amin(-1))
    got = pb.sum(xd, axis=-1).cpu()
    if dt.is_floating_point:
        tol = {torch.float32: 1e-5, torch.bfloat16: 1e-2, torch.float64: 1e-12}[dt]
        np.testing.assert_allclose(got.double().numpy(), x.double().sum(-1).numpy(), rtol=tol,
                                   atol=tol * float(x.double().abs().sum(-1).max()))
    else:
        assert torch.equal(got.long(), x.long().sum(-1) % 256 if dt == torch.uint8 else x.long().sum(-1))
    if dt.is_floating_point:
        m = pb.mean(xd, axis=-1).double().cpu().numpy()
        np.testing.assert_allclose(m, x.double().mean(-1).numpy(), rtol=1e-2 if dt == torch.bfloat16 else 1e-5,
                                   atol=1e-2 if dt == torch.bfloat16 else 1e-5)
