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
