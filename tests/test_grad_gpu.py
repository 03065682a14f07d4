"""Backward of sum / mean / add / subtract / multiply against the
oracle: the reference computes them with the same two engines (reduce_grad.h:30-85,
elementwise_grad.h:29-105), so the oracle's reduce/elementwise are the checkers."""
import numpy as np
import pytest
import torch

import cases
import gpu_util as G
from paddle2_b200 import grad

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,axis,keepdim", [((5, 6, 10), [0], False), ((5, 6, 10), [-1], True),
                                                ((5, 6, 10), None, False), ((2, 4, 3, 5, 6, 10), [2, 4], False),
                                                ((33, 1025), [1], False), ((100, 1, 1), [1, 2], True)])
def test_sum_and_mean_grad(oracle, shape, axis, keepdim):
    rng = np.random.default_rng(1)
    x = rng.random(shape).astype(np.float32)
    out = oracle.reduce("sum", x, axis, keepdim)
    dout = rng.random(out.shape).astype(np.float32)
    axes, reduce_all = oracle._axes_list(axis, x.ndim)
    kd = [1 if (reduce_all or i in axes) else s for i, s in enumerate(shape)]
    want = np.broadcast_to(dout.reshape(kd), shape)
    got = G.to_np(grad.sum_grad(G.to_dev(x), G.to_dev(dout), axis, keepdim))
    assert np.array_equal(got, want)
    n = x.size // max(out.size, 1)
    gotm = G.to_np(grad.mean_grad(G.to_dev(x), G.to_dev(dout), axis, keepdim))
    np.testing.assert_allclose(gotm, want * np.float32(1.0 / n), rtol=1e-6)


@pytest.mark.parametrize("i", range(len(cases.ADD_CASES)))
def test_add_sub_mul_grad(oracle, i):
    xs, ys, axis, src = cases.ADD_CASES[i]
    rng = np.random.default_rng(50 + i)
    x = cases.make_input(rng, "float32", xs, "signed")
    y = cases.make_input(rng, "float32", ys, "signed")
    out_shape, xp, yp = oracle.broadcast_out_shape(xs, ys, axis)
    dout = rng.random(out_shape).astype(np.float32)

    def back(g, padded, shape):   # sum over the dims where the operand was broadcast
        dims = [k for k, (a, o) in enumerate(zip(padded, out_shape)) if a == 1 and o != 1]
        r = oracle.reduce("sum", g, dims, False) if dims else g
        return np.asarray(r).reshape(shape)

    dx, dy = grad.add_grad(G.to_dev(x), G.to_dev(y), G.to_dev(dout), axis)
    G.assert_reduce_close(G.to_np(dx), back(dout, xp, xs), dout, None, "prod", 1e-5)
    G.assert_reduce_close(G.to_np(dy), back(dout, yp, ys), dout, None, "prod", 1e-5)
    dx, dy = grad.subtract_grad(G.to_dev(x), G.to_dev(y), G.to_dev(dout), axis)
    np.testing.assert_allclose(G.to_np(dy), -back(dout, yp, ys), rtol=1e-5, atol=1e-6)
    gx = dout * y.reshape(yp)
    gy = dout * x.reshape(xp)
    dx, dy = grad.multiply_grad(G.to_dev(x), G.to_dev(y), G.to_dev(dout), axis)
    np.testing.assert_allclose(G.to_np(dx), back(np.ascontiguousarray(gx), xp, xs), rtol=1e-4, atol=1e-5, err_msg=src)
    np.testing.assert_allclose(G.to_np(dy), back(np.ascontiguousarray(gy), yp, ys), rtol=1e-4, atol=1e-5, err_msg=src)
