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


def _keepdims_shape(oracle, shape, axis):
    axes, reduce_all = oracle._axes_list(axis, len(shape))
    return [1 if (reduce_all or i in axes) else s for i, s in enumerate(shape)], \
        (list(range(len(shape))) if reduce_all else list(axes))


@pytest.mark.parametrize("dt", ["float32", "float64", "int32", "int64"])
@pytest.mark.parametrize("shape,axis,keepdim", [((5, 6, 10), [-1], False), ((5, 6, 10), [0], True),
                                                ((5, 6, 10), None, False), ((2, 4, 3, 5, 6, 10), [2, 4], False),
                                                ((33, 1025), [1], False), ((300, 1024), [0], False)])
def test_max_min_grad(oracle, dt, shape, axis, keepdim):
    """ReduceMax/MinGradKernel (gpu/reduce_kernel.cu:74-119,144-187): dout * T(Equal(x, out)),
    ties included -- checked against the oracle's Equal and Multiply functors."""
    rng = np.random.default_rng(8)
    x = cases.map_operand(rng, dt, shape, "ties")       # few distinct values: every row has ties
    kd, _ = _keepdims_shape(oracle, shape, axis)
    for op, fn in (("max", grad.max_grad), ("min", grad.min_grad)):
        out = oracle.reduce(op, x, axis, keepdim)
        dout = cases.map_operand(rng, dt, out.shape)
        want = oracle.map("equal_mul", [x, np.asarray(out).reshape(kd), np.asarray(dout).reshape(kd)])
        got = G.to_np(fn(G.to_dev(x), G.to_dev(out), G.to_dev(dout), axis, keepdim))
        assert got.shape == x.shape
        assert np.array_equal(got, want), (op, dt, shape, axis)
        # every tie carries the gradient: sum(dx) over the reduced dims = dout * (#ties)
        if dt == "float64":
            ties = (x == np.asarray(out).reshape(kd)).sum(axis=tuple(_keepdims_shape(oracle, shape, axis)[1]),
                                                          keepdims=True)
            np.testing.assert_allclose(got.sum(axis=tuple(_keepdims_shape(oracle, shape, axis)[1]), keepdims=True),
                                       np.asarray(dout).reshape(kd) * ties, rtol=1e-12)


@pytest.mark.parametrize("shape,axis", [((5, 6, 10), [-1]), ((5, 6, 10), [0, 2]), ((64, 513), None),
                                        ((40, 1024), [0])])
def test_amax_amin_grad(oracle, shape, axis):
    """ReduceCudaAMaxAMinGrad (gpu/reduce_amin_amax_common.h:28-107): the gradient is shared by the
    ties; the oracle composes the reference's own four steps (Equal, Sum, Multiply, Divide)."""
    rng = np.random.default_rng(9)
    x = cases.map_operand(rng, "float32", shape, "ties")
    kd, dims = _keepdims_shape(oracle, shape, axis)
    for op, fn in (("max", grad.amax_grad), ("min", grad.amin_grad)):
        out = np.asarray(oracle.reduce(op, x, axis, False)).reshape(kd)
        dout = cases.map_operand(rng, "float32", kd)
        eq = oracle.map("equal_t", [out, x])
        count = np.asarray(oracle.reduce("sum", eq, dims, True)).reshape(kd)
        want = oracle.elementwise("divide", [oracle.elementwise("multiply", [dout, eq]), count])
        got = G.to_np(fn(G.to_dev(x), G.to_dev(out.reshape(np.asarray(oracle.reduce(op, x, axis, False)).shape)),
                         G.to_dev(dout.reshape(np.asarray(oracle.reduce(op, x, axis, False)).shape)), axis, False))
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=0)
        np.testing.assert_allclose(got.sum(axis=tuple(dims), keepdims=True), dout, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("i", range(len(cases.ADD_CASES)))
def test_divide_maximum_minimum_grad(oracle, i):
    xs, ys, axis, src = cases.ADD_CASES[i]
    rng = np.random.default_rng(90 + i)
    x = cases.map_operand(rng, "float32", xs, "ties")
    y = cases.map_operand(rng, "float32", ys, "ties")
    out_shape, xp, yp = oracle.broadcast_out_shape(xs, ys, axis)
    dout = cases.map_operand(rng, "float32", out_shape)
    xb, yb = x.reshape(xp), y.reshape(yp)

    def back(g, padded, shape):
        dims = [k for k, (a, o) in enumerate(zip(padded, out_shape)) if a == 1 and o != 1]
        r = oracle.reduce("sum", np.ascontiguousarray(g), dims, False) if dims else g
        return np.asarray(r).reshape(shape)

    for fn, nx, ny in ((grad.maximum_grad, "max_grad_x", "max_grad_y"),
                       (grad.minimum_grad, "min_grad_x", "min_grad_y")):
        dx, dy = fn(G.to_dev(x), G.to_dev(y), G.to_dev(dout), axis)
        wx = back(oracle.map(nx, [xb, yb, dout]), xp, xs)
        wy = back(oracle.map(ny, [xb, yb, dout]), yp, ys)
        np.testing.assert_allclose(G.to_np(dx), wx, rtol=1e-4, atol=1e-4, err_msg=src)
        np.testing.assert_allclose(G.to_np(dy), wy, rtol=1e-4, atol=1e-4, err_msg=src)
        # the two masks partition the output: dx + dy accounts for all of dout
        np.testing.assert_allclose(float(G.to_np(dx).astype(np.float64).sum() + G.to_np(dy).astype(np.float64).sum()),
                                   float(dout.astype(np.float64).sum()), rtol=1e-4, atol=1e-3)
    yq = (np.abs(y) + 0.5).astype(np.float32)
    out = oracle.elementwise("divide", [x, yq], axis=axis)
    dx, dy = grad.divide_grad(G.to_dev(x), G.to_dev(yq), G.to_dev(out), G.to_dev(dout), axis)
    yqb = yq.reshape(yp)
    wx = back(oracle.elementwise("divide", [dout, yqb]), xp, xs)
    wy = back(oracle.map("div_grad_y", [dout, out, yqb]), yp, ys)
    np.testing.assert_allclose(G.to_np(dx), wx, rtol=1e-4, atol=1e-4, err_msg=src)
    np.testing.assert_allclose(G.to_np(dy), wy, rtol=1e-4, atol=1e-4, err_msg=src)
