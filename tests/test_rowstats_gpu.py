"""Row mean/variance (C4) and paddle.var parity."""
import numpy as np
import pytest
import torch

import gpu_util as G
import paddle2_b200 as pb

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,cols", [(4096, 1024), (300, 8192), (64, 1000), (17, 100003), (5000, 37), (2, 1)])
def test_fp32_rows(oracle, rows, cols):
    x = np.random.default_rng(rows + cols).random((rows, cols)).astype(np.float32)
    mean, var = oracle.rowstats_f32(x, unbiased=True)
    m, v = pb.row_mean_var(G.to_dev(x))
    np.testing.assert_allclose(G.to_np(m), mean, rtol=G.FP32_RTOL)
    np.testing.assert_allclose(G.to_np(v), var, rtol=1e-4, atol=1e-7)
    v2 = pb.var(G.to_dev(x), axis=-1)
    assert torch.equal(v2, v)
    vb = pb.var(G.to_dev(x), axis=-1, unbiased=False, keepdim=True)
    np.testing.assert_allclose(G.to_np(vb)[:, 0], np.var(x.astype(np.float64), axis=-1), rtol=1e-4, atol=1e-7)


@pytest.mark.parametrize("rows,cols", [(2048, 8192), (4096, 2048), (33, 1001)])
def test_bf16_rows(oracle, rows, cols):
    rng = np.random.default_rng(3)
    x32 = G.bf16_round((rng.standard_normal((rows, cols)) * np.sqrt(2) + 0.5).astype(np.float32))
    mean, var = oracle.rowstats_f32(x32, unbiased=True)
    xt = torch.from_numpy(x32).cuda().bfloat16()
    m, v = pb.row_mean_var(xt)
    assert m.dtype == torch.bfloat16
    np.testing.assert_allclose(G.to_np(m), mean, rtol=G.BF16_RTOL, atol=1e-3)
    np.testing.assert_allclose(G.to_np(v), var, rtol=G.BF16_RTOL)
    m32, v32 = pb.row_mean_var(xt, out_dtype="float32")   # fp32 outputs for a tighter check
    np.testing.assert_allclose(G.to_np(m32), mean, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(G.to_np(v32), var, rtol=1e-4)


def test_var_other_axes_uses_the_composite(oracle):
    x = np.random.default_rng(4).random((6, 50, 33)).astype(np.float32)
    for axis in (0, [0, 2], None, 1):
        want = oracle.var_composite(x, axis=axis)
        got = G.to_np(pb.var(G.to_dev(x), axis=axis))
        np.testing.assert_allclose(got, want, rtol=1e-4)
    x64 = x.astype(np.float64)
    np.testing.assert_allclose(G.to_np(pb.var(G.to_dev(x64), axis=-1)), np.var(x64, axis=-1, ddof=1), rtol=1e-10)


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("cols", [8192, 4096, 3000])
def test_16bit_rows_poorly_conditioned(tdt, cols):
    """Rows whose first vector sits far from the rest, rows with a large mean and a small spread,
    constant rows: the cases a one-sweep (shifted sum of squares) formulation loses digits on.
    The kernel is the exact two-sweep form held in registers, so they must be as accurate as any
    other row; fp32 outputs, float64 numpy as the judge.  (One-sweep variants were measured and
    dropped: see DESIGN.md, row statistics.)"""
    rng = np.random.default_rng(cols)
    rows = 512
    x = rng.standard_normal((rows, cols)).astype(np.float32)
    x[0::4, :8] = 1000.0                    # every element of the first vector is a 1000-sigma outlier
    x[1::4, :8] = -300.0
    x[2::4, :] = x[2::4, :] * 0.25 + 200.0  # large mean, small spread (K is near the mean: stays one-sweep)
    x[3::4, 3] = 500.0                      # a single outlier inside the first vector
    xt = torch.from_numpy(x).to(tdt)
    x64 = xt.double().numpy()
    m, v = pb.row_mean_var(xt.cuda(), out_dtype="float32")
    np.testing.assert_allclose(G.to_np(m), x64.mean(axis=1), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(G.to_np(v), x64.var(axis=1, ddof=1), rtol=1e-4)
    # constant rows: variance exactly 0, mean exact
    c = torch.full((64, cols), 3.5, dtype=tdt)
    m, v = pb.row_mean_var(c.cuda(), out_dtype="float32")
    assert torch.all(v == 0) and torch.all(m == 3.5)


@pytest.mark.parametrize("tdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,cols", [(33, 32000), (7, 50257), (5, 100003), (64, 16384)])
def test_long_rows_one_sweep(tdt, rows, cols):
    """rows that fit no register kernel go through the one-sweep pairwise-moments kernel (any
    alignment); it must be as accurate as the two-sweep form, including on rows with a large
    mean and a small spread, where a sum / sum-of-squares formulation would cancel"""
    gen = torch.Generator().manual_seed(cols)
    x = torch.randn((rows, cols), generator=gen)
    x[1] = x[1] * 0.01 + 300.0
    x[2, :16] = 5000.0
    xt = x.to(tdt)
    x64 = xt.double()
    m, v = pb.row_mean_var(xt.cuda(), out_dtype="float32") if tdt != torch.float32 else pb.row_mean_var(xt.cuda())
    np.testing.assert_allclose(m.cpu().numpy(), x64.mean(-1).numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(v.cpu().numpy(), x64.var(-1, unbiased=True).numpy(), rtol=1e-4)


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("rows,cols", [(64, 301), (333, 501), (129, 701), (65, 1001), (64, 1025), (777, 1027), (130, 2049), (65, 2051), (100, 3001), (67, 4097), (3, 1025),
                                       (70, 6001), (66, 8191), (64, 8193)])
def test_unaligned_medium_rows_warp_peel(rows, cols, dt):
    """1..16 KiB rows that are not whole aligned vectors: a warp per row with the row in registers
    (head / vectors / tail); also from a base one element past a 16-byte boundary"""
    import numpy as np
    import paddle2_b200 as pb
    gen = torch.Generator().manual_seed(rows + 31 * cols)
    flat = (torch.randn(rows * cols + 1, generator=gen) * 2 + 0.5).to(dt)
    fd = flat.cuda()
    for off in (0, 1):
        x = flat[off:off + rows * cols].view(rows, cols)
        xd = fd[off:off + rows * cols].view(rows, cols)
        m, v = pb.row_mean_var(xd, out_dtype="float32")
        x64 = x.double()
        np.testing.assert_allclose(m.cpu().numpy(), x64.mean(-1).numpy(), rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(v.cpu().numpy(), x64.var(-1, unbiased=True).numpy(), rtol=1e-4)
