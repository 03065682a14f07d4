"""Backward of the hot-path ops, composed from the same two engines
exactly as the reference composes them:

  sum_grad / mean_grad   broadcast of out_grad through the broadcast engine
                         (paddle/phi/kernels/gpu/reduce_grad.h:30-85, gpu/reduce_kernel.cu:42-232)
  add/subtract/multiply_grad   elementwise op, then phi::SumKernel over the broadcast dims
                         (paddle/phi/kernels/gpu/elementwise_grad.h:29-105,173-304;
                          funcs::GetReduceDim, elementwise_base.h:445+)

Every value is computed by libb200phi.so (b200phi_elementwise / b200phi_reduce)."""
import torch

from . import _lib as L
from . import ops


def expand(x, shape):
    """x broadcast (numpy rules) into `shape`: the unary BroadcastKernel call of ReduceGrad."""
    return ops._elementwise(L.ADD, [x], out_shape=list(shape))


def _keepdim_view(out_grad, x_shape, axis):
    rank = len(x_shape)
    axes, reduce_all = ops._norm_axes(axis, rank)
    dims = [1 if (reduce_all or i in axes) else s for i, s in enumerate(x_shape)]
    n = 1
    for i, s in enumerate(x_shape):
        if reduce_all or i in axes:
            n *= s
    return out_grad.contiguous().reshape(dims), n


def sum_grad(x, out_grad, axis=None, keepdim=False):
    """d/dx of paddle.sum: out_grad, reshaped with the reduced dims as 1, broadcast to x's shape
    (ReduceGradKernel, reduce_grad.h:44-82)."""
    g, _ = _keepdim_view(out_grad, list(x.shape), axis)
    if g.dtype != x.dtype:
        g = ops.cast(g, x.dtype)
    return expand(g, x.shape)


def mean_grad(x, out_grad, axis=None, keepdim=False):
    """d/dx of paddle.mean: broadcast(out_grad) / reduce_num, one fused kernel
    (kps::DivideFunctor in ReduceMeanGradKernel, gpu/reduce_kernel.cu)."""
    g, n = _keepdim_view(out_grad, list(x.shape), axis)
    return ops._elementwise(L.ADD, [g], act=L.ACT_SCALE, scale_v=1.0 / n, out_shape=list(x.shape))


def reduce_dims_for(dst_shape, src_shape, axis=-1):
    """funcs::GetReduceDim(dst_dims, src_dims, axis): the dims of `src` over which a gradient
    must be summed to get back to the (broadcast) operand shape `dst`."""
    dst_shape, src_shape = list(dst_shape), list(src_shape)
    ax = len(src_shape) - len(dst_shape) if axis < 0 else axis
    dims = []
    for i in range(len(src_shape)):
        if i < ax or i >= ax + len(dst_shape):
            dims.append(i)
        elif dst_shape[i - ax] == 1 and src_shape[i] != 1:
            dims.append(i)
    return dims


def _reduce_to(g, shape, axis=-1):
    if list(g.shape) == list(shape):
        return g
    dims = reduce_dims_for(shape, g.shape, axis)
    r = ops.sum(g, axis=dims, keepdim=False, dtype=g.dtype) if dims else g
    return r.reshape(list(shape))


def add_grad(x, y, out_grad, axis=-1):
    """(dx, dy) of x + y with broadcasting (AddGradFunc, elementwise_grad.h)."""
    y_axis = axis if len(y.shape) <= len(x.shape) else -1
    x_axis = axis if len(x.shape) < len(y.shape) else -1
    return _reduce_to(out_grad, x.shape, x_axis), _reduce_to(out_grad, y.shape, y_axis)


def subtract_grad(x, y, out_grad, axis=-1):
    dx, dy = add_grad(x, y, out_grad, axis)
    return dx, ops.scale(dy, -1.0)


def multiply_grad(x, y, out_grad, axis=-1):
    """dx = sum(out_grad * y), dy = sum(out_grad * x) over the broadcast dims."""
    gx = ops.multiply(out_grad, y, axis=axis if len(y.shape) < len(out_grad.shape) else -1)
    gy = ops.multiply(out_grad, x, axis=axis if len(x.shape) < len(out_grad.shape) else -1)
    y_axis = axis if len(y.shape) <= len(x.shape) else -1
    x_axis = axis if len(x.shape) < len(y.shape) else -1
    return _reduce_to(gx, x.shape, x_axis), _reduce_to(gy, y.shape, y_axis)


def _minmax_grad(x, out, out_grad, axis):
    o, _ = _keepdim_view(out, list(x.shape), axis)
    g, _ = _keepdim_view(out_grad, list(x.shape), axis)
    return ops._map("equal_mul", [x, o, g])


def max_grad(x, out, out_grad, axis=None, keepdim=False):
    """d/dx of paddle.max: out_grad * T(Equal(x, out)) with out and out_grad broadcast over the
    reduced dims.  The reference runs BroadcastKernel(EqualFunctor) into a temporary, then
    MultiplyKernel (ReduceMaxGradKernel, gpu/reduce_kernel.cu:144-187); here it is one ternary
    launch, no temporary.  Every position that ties with the max receives the full gradient."""
    return _minmax_grad(x, out, out_grad, axis)


def min_grad(x, out, out_grad, axis=None, keepdim=False):
    """ReduceMinGradKernel, gpu/reduce_kernel.cu:74-119: same form as max_grad."""
    return _minmax_grad(x, out, out_grad, axis)


def _aminmax_grad(x, out, out_grad, axis):
    o, _ = _keepdim_view(out, list(x.shape), axis)
    g, _ = _keepdim_view(out_grad, list(x.shape), axis)
    rank = len(x.shape)
    axes, reduce_all = ops._norm_axes(axis, rank)
    dims = list(range(rank)) if reduce_all else list(axes)
    mask = ops._map("equal_t", [x, o])
    count = ops.sum(mask, axis=dims, keepdim=True, dtype=x.dtype) if dims else mask
    share = ops.divide(g, count.reshape(list(g.shape)))
    return ops._map("equal_mul", [x, o, share])


def amax_grad(x, out, out_grad, axis=None, keepdim=False):
    """d/dx of paddle.amax: the gradient is shared evenly between the positions that tie with
    the max: (out_grad * eq) / sum(eq) (ReduceCudaAMaxAMinGrad,
    gpu/reduce_amin_amax_common.h:28-107).  out_grad / sum(eq) is formed on the small tensor
    first, which gives the same quotient at the ties and 0 elsewhere."""
    return _aminmax_grad(x, out, out_grad, axis)


def amin_grad(x, out, out_grad, axis=None, keepdim=False):
    return _aminmax_grad(x, out, out_grad, axis)


def _bcast_axis(a, b, axis):
    return axis if len(b.shape) < len(a.shape) else -1


def divide_grad(x, y, out, out_grad, axis=-1):
    """dx = sum(dout / y), dy = sum(-dout * ((out / y) / y)) over the broadcast dims
    (DivGradXFunctor / DivGradYFunctor, funcs/elementwise_functor.h:157-179;
    ElementwiseDivGrad, gpu/elementwise_grad.h)."""
    gx = ops.divide(out_grad, y, axis=_bcast_axis(out_grad, y, axis))
    gy = ops._map("div_grad_y", [out_grad, out, _align(y, out.shape, x.shape, axis)])
    y_axis = axis if len(y.shape) <= len(x.shape) else -1
    x_axis = axis if len(x.shape) < len(y.shape) else -1
    return _reduce_to(gx, x.shape, x_axis), _reduce_to(gy, y.shape, y_axis)


def _align(t, out_shape, other_shape, axis):
    """View the lower-rank operand with 1s so that numpy-style (right-aligned) broadcasting
    places it where the legacy `axis` attribute would (common_shape.h:35-92)."""
    r = len(out_shape)
    if len(t.shape) == r:
        return t
    off = r - len(t.shape) if axis < 0 else axis
    dims = [1] * off + list(t.shape) + [1] * (r - off - len(t.shape))
    return t.contiguous().reshape(dims)


def _select_grad(name_x, name_y, x, y, out_grad, axis):
    shape = list(out_grad.shape)
    xa, ya = _align(x, shape, y.shape, axis), _align(y, shape, x.shape, axis)
    gx = ops._map(name_x, [xa, ya, out_grad])
    gy = ops._map(name_y, [xa, ya, out_grad])
    y_axis = axis if len(y.shape) <= len(x.shape) else -1
    x_axis = axis if len(x.shape) < len(y.shape) else -1
    return _reduce_to(gx, x.shape, x_axis), _reduce_to(gy, y.shape, y_axis)


def maximum_grad(x, y, out_grad, axis=-1):
    """dx = dout * (x > y), dy = dout * (x <= y), summed over the broadcast dims
    (MaxGradXFunctor / MaxGradYFunctor, funcs/elementwise_functor.h:478-490)."""
    return _select_grad("max_grad_x", "max_grad_y", x, y, out_grad, axis)


def minimum_grad(x, y, out_grad, axis=-1):
    """dx = dout * (x < y), dy = dout * (x >= y) (MinGradXFunctor / MinGradYFunctor, :513-524)."""
    return _select_grad("min_grad_x", "min_grad_y", x, y, out_grad, axis)
