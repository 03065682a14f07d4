"""Reduce-side plug-ins: squared_l2_norm, p_norm, mean_all, reduce_as, dist,
broadcast_tensors, equal_all and logsumexp against the oracle and the reference-made golden vectors."""
import os

import numpy as np
import pytest
import torch

import cases
import gpu_util as G
import paddle2_b200 as pb
from paddle2_b200 import ops

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "phi_ref_golden.npz"))


@pytest.mark.parametrize("i", range(len(cases.PNORM_CASES)))
def test_p_norm_golden(golden, i):
    shape, axis, porder, keepdim, asvector, src = cases.PNORM_CASES[i]
    x = golden[f"pnorm_{i}_x"]
    got = G.to_np(ops.p_norm(G.to_dev(x), porder, axis, keepdim=keepdim, asvector=asvector))
    want = golden[f"pnorm_{i}_y"]
    assert got.shape == want.shape, src
    if porder in (0.0, float("inf"), float("-inf")):
        assert np.array_equal(got, want), src
    else:
        np.testing.assert_allclose(got, want, rtol=1e-5, err_msg=src)


@pytest.mark.parametrize("dt,rtol", [("float32", 1e-5), ("float64", 1e-12)])
@pytest.mark.parametrize("shape,axis", [((64, 4096), -1), ((4096, 64), 0), ((7, 33, 129), 1), ((100003,), 0)])
def test_norm_ops_shapes(oracle, dt, rtol, shape, axis):
    """every kernel kind (row, column, generic, split) carries the transform"""
    x = cases.map_operand(np.random.default_rng(5), dt, shape)
    x.reshape(-1)[::5] = 0
    dx = G.to_dev(x)
    for porder in (0.0, 1.0, 2.0, float("inf"), float("-inf"), 3.0):
        got = G.to_np(ops.p_norm(dx, porder, axis))
        want = oracle.p_norm(x, porder, axis)
        assert got.shape == want.shape
        if porder in (0.0, float("inf"), float("-inf")):
            assert np.array_equal(got, want), (porder, shape)
        else:
            np.testing.assert_allclose(got, want, rtol=max(rtol, 1e-11) * 20, err_msg=f"p={porder} {shape}")
    got = G.to_np(ops.squared_l2_norm(dx))
    assert got.shape == (1,)
    np.testing.assert_allclose(got, (x.astype(np.float64) ** 2).sum(), rtol=max(rtol, 1e-11) * 20)
    np.testing.assert_allclose(G.to_np(ops.mean_all(dx)), x.astype(np.float64).mean(), rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
def test_norm_ops_16bit(tdt):
    x = (torch.randn((257, 2048), generator=torch.Generator().manual_seed(1)) * 2).to(tdt)
    x64 = x.double().numpy()
    for porder, want in ((1.0, np.abs(x64).sum(-1)), (2.0, np.sqrt((x64 ** 2).sum(-1))),
                         (float("inf"), np.abs(x64).max(-1)), (0.0, (x64 != 0).sum(-1))):
        got = ops.p_norm(x.cuda(), porder, -1)
        assert got.dtype == tdt
        np.testing.assert_allclose(got.double().cpu().numpy(), want, rtol=1e-2)
    got = ops.squared_l2_norm(x[:4].cuda())   # small enough for an fp16 result
    np.testing.assert_allclose(got.double().cpu().numpy(), (x64[:4] ** 2).sum(), rtol=1e-2)


def test_squared_l2_norm_golden(golden):
    got = G.to_np(ops.squared_l2_norm(G.to_dev(golden["sql2_x"])))
    np.testing.assert_allclose(got, golden["sql2_y"], rtol=1e-5)


@pytest.mark.parametrize("i", range(len(cases.LOGSUMEXP_CASES)))
def test_logsumexp_golden(oracle, golden, i):
    shape, axis, keepdim, src = cases.LOGSUMEXP_CASES[i]
    x = golden[f"lse_{i}_x"]
    got = G.to_np(ops.logsumexp(G.to_dev(x), axis if axis else None, keepdim))
    want = golden[f"lse_{i}_y"] if f"lse_{i}_y" in golden.files else oracle.logsumexp(x, axis, keepdim)
    assert got.shape == want.shape, src
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6, err_msg=src)


@pytest.mark.parametrize("dt,rtol", [("float32", 1e-5), ("float64", 1e-12)])
@pytest.mark.parametrize("rows,cols", [(64, 32000), (1000, 10), (33, 1), (3, 100003), (300, 8192), (17, 4099)])
def test_logsumexp_rows(dt, rtol, rows, cols):
    rng = np.random.default_rng(rows + cols)
    x = (rng.standard_normal((rows, cols)) * 4).astype(dt)
    x[0, : min(cols, 5)] = 80.0          # exp overflows in fp32 without the max shift
    x[-1, :] = -90.0                     # ... and underflows
    x64 = x.astype(np.float64)
    m = x64.max(-1, keepdims=True)
    want = (m + np.log(np.exp(x64 - m).sum(-1, keepdims=True)))[:, 0]
    got = G.to_np(ops.logsumexp(G.to_dev(x), -1))
    np.testing.assert_allclose(got, want, rtol=max(rtol, 1e-11) * 10, atol=1e-6)
    # shift invariance: lse(x + c) = lse(x) + c
    got2 = G.to_np(ops.logsumexp(G.to_dev((x64 + 3.0).astype(dt)), -1))
    np.testing.assert_allclose(got2, want + 3.0, rtol=max(rtol, 1e-11) * 10, atol=1e-5)


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
def test_logsumexp_16bit_and_special_values(tdt):
    x = (torch.randn((129, 32000), generator=torch.Generator().manual_seed(2)) * 3).to(tdt)
    got = ops.logsumexp(x.cuda(), -1)
    assert got.dtype == tdt
    want = torch.logsumexp(x.double(), -1).numpy()
    np.testing.assert_allclose(got.double().cpu().numpy(), want, rtol=1e-2)
    # rows of -inf give -inf; a -inf entry is ignored
    y = torch.full((4, 1000), float("-inf"))
    y[1, 7] = 2.0
    y[2, :] = torch.linspace(-5, 5, 1000)
    got = ops.logsumexp(y.cuda(), -1).cpu()
    assert got[0] == float("-inf") and got[3] == float("-inf")
    assert abs(float(got[1]) - 2.0) < 1e-6
    np.testing.assert_allclose(float(got[2]), float(torch.logsumexp(y[2].double(), -1)), rtol=1e-6)


def test_reduce_as_dist_broadcast_equal_all(oracle):
    rng = np.random.default_rng(12)
    x = rng.standard_normal((6, 5, 40)).astype(np.float32)
    for tshape in [(5, 40), (5, 1), (1, 40), (6, 1, 40), (6, 5, 40), (40,)]:
        t = np.zeros(tshape, np.float32)
        got = G.to_np(ops.reduce_as(G.to_dev(x), G.to_dev(t)))
        off = x.ndim - len(tshape)
        dims = tuple(list(range(off)) + [i + off for i, s in enumerate(tshape) if s == 1 and x.shape[i + off] != 1])
        want = x.astype(np.float64).sum(axis=dims).reshape(tshape) if dims else x
        assert got.shape == tuple(tshape)
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5)
    y = rng.standard_normal((5, 1)).astype(np.float32)
    for p in (0.0, 1.0, 2.0, float("inf"), float("-inf"), 3.0):
        got = float(G.to_np(ops.dist(G.to_dev(x), G.to_dev(y), p)))
        d = np.abs(x.astype(np.float64) - y.astype(np.float64)).reshape(-1)
        want = {0.0: float((d != 0).sum()), float("inf"): d.max(), float("-inf"): d.min()}.get(
            p, float((d ** p).sum() ** (1.0 / p)) if p not in (0.0,) and np.isfinite(p) else None)
        np.testing.assert_allclose(got, want, rtol=1e-5)
    a, b, c = (rng.standard_normal(s).astype(np.float32) for s in [(1, 64, 112), (1, 64, 1), (3, 1, 112)])
    outs = ops.broadcast_tensors([G.to_dev(a), G.to_dev(b), G.to_dev(c)])
    for o, src in zip(outs, (a, b, c)):
        assert np.array_equal(G.to_np(o), np.broadcast_to(src, (3, 64, 112)))
    xe = G.to_dev(x)
    assert bool(ops.equal_all(xe, xe.clone())) is True
    x2 = x.copy()
    x2[3, 2, 1] += 1
    assert bool(ops.equal_all(xe, G.to_dev(x2))) is False
    assert bool(ops.equal_all(xe, G.to_dev(x[:5]))) is False


def test_errors():
    from paddle2_b200._lib import B200PhiError
    with pytest.raises(B200PhiError):
        ops.p_norm(torch.ones((4, 4), dtype=torch.int32).cuda(), 2.0, -1)
    with pytest.raises(B200PhiError):
        ops.logsumexp(torch.ones((0, 4)).cuda(), -1)
    with pytest.raises(B200PhiError):
        ops.logsumexp(torch.ones((4, 4), dtype=torch.int32).cuda(), -1)
