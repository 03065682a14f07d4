"""paddle.* look-alikes for the hot path, on torch CUDA tensors, through the C ABI.

Mirrors the reference's Python entry points (argument names, defaults, dtype and
shape rules) so the parity tests read like the reference's own:

  sum   python/paddle/tensor/math.py:1475-1625     add   math.py:682
  max   math.py:2870   min math.py:3029   prod math.py:4779   any math.py:5119
  mean  python/paddle/tensor/stat.py:45            var   stat.py:141-198
  argmax python/paddle/tensor/search.py:175

torch is only the owner of device memory and streams here (torch.empty,
torch.cuda.current_stream); every value is computed by libb200phi.so.  Shape and
dtype inference that Paddle runs before the kernel (InferMeta) is re-stated in
_reduce_out_shape / _sum_out_dtype / broadcast_shape with citations.
"""
import ctypes as C

import torch

from . import _lib as L

_T2P = {
    torch.bool: L.BOOL,
    torch.uint8: L.UINT8,
    torch.int8: L.INT8,
    torch.int16: L.INT16,
    torch.int32: L.INT32,
    torch.int64: L.INT64,
    torch.float32: L.FLOAT32,
    torch.float64: L.FLOAT64,
    torch.float16: L.FLOAT16,
    torch.bfloat16: L.BFLOAT16,
    torch.complex64: L.COMPLEX64,      # sum / mean / add / subtract
    torch.complex128: L.COMPLEX128,
    torch.float8_e4m3fn: L.FLOAT8_E4M3FN,  # max / amax
    torch.float8_e5m2: L.FLOAT8_E5M2,
}
_NAME2T = {
    "bool": torch.bool, "uint8": torch.uint8, "int8": torch.int8, "int16": torch.int16,
    "int32": torch.int32, "int64": torch.int64, "float32": torch.float32,
    "float64": torch.float64, "float16": torch.float16, "bfloat16": torch.bfloat16,
    "complex64": torch.complex64, "complex128": torch.complex128,
}


def _to_torch_dtype(dtype):
    if dtype is None or isinstance(dtype, torch.dtype):
        return dtype
    if isinstance(dtype, str) and dtype in _NAME2T:
        return _NAME2T[dtype]
    raise TypeError(f"unsupported dtype {dtype!r}")


def _phi_dtype(t):
    try:
        return _T2P[t]
    except KeyError:
        raise TypeError(f"dtype {t} is not supported by the b200phi path") from None


def _require_cuda(*tensors):
    for t in tensors:
        if not isinstance(t, torch.Tensor):
            raise TypeError("expected a torch.Tensor holding device memory")
        if not t.is_cuda:
            raise RuntimeError(
                "b200phi has no CPU path: tensors must live on a CUDA device "
                "(the reference's CPU kernels are only used as the test oracle)")


def _stream_ptr(x):
    return C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)


def _norm_axes(axis, rank):
    """axis argument -> (sorted unique non-negative axes, reduce_all).
    ReduceInferDim, paddle/phi/infermeta/unary.cc:3631-3700."""
    if axis is None:
        return list(range(rank)), True
    if isinstance(axis, int):
        axis = [axis]
    axis = list(axis)
    if len(axis) == 0:
        return list(range(rank)), True
    lo, hi = (-rank, rank) if rank > 0 else (-1, 1)
    out = set()
    for a in axis:
        if not (lo <= a < hi):
            raise ValueError(
                f"The reduce dim index should be in the range [-dimension(X), dimension(X)) "
                f"which dimension = {rank}, but received dim index = {a}.")
        out.add(a + rank if a < 0 else a)
    if rank == 0:
        return [], True
    return sorted(out), len(out) == rank


def _reduce_out_shape(shape, axes, reduce_all, keepdim):
    rank = len(shape)
    if reduce_all:
        return [1] * rank if keepdim else []
    if keepdim:
        return [1 if i in axes else s for i, s in enumerate(shape)]
    return [s for i, s in enumerate(shape) if i not in axes]


def _reduce(op, x, axis, keepdim, out_dtype, out=None):
    _require_cuda(x)
    x = x.contiguous()  # PrepareData/Trans2Contiguous does this before phi kernels
    rank = x.dim()
    axes, reduce_all = _norm_axes(axis, rank)
    out_shape = _reduce_out_shape(list(x.shape), axes, reduce_all, keepdim)
    if out is None:
        out = torch.empty(out_shape, dtype=out_dtype, device=x.device)
    else:  # caller-provided result storage (the phi contract: `DenseTensor* out`)
        _require_cuda(out)
        if list(out.shape) != out_shape or out.dtype != out_dtype or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous {out_dtype} tensor of shape {out_shape}")
    d = L.ReduceDesc()
    d.op = op
    d.in_dtype = _phi_dtype(x.dtype)
    d.out_dtype = _phi_dtype(out_dtype)
    d.rank = rank
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.naxes = 0 if reduce_all else len(axes)
    for i, a in enumerate(axes if not reduce_all else []):
        d.axes[i] = a
    d.reduce_all = 1 if reduce_all else 0
    d.x = x.data_ptr()
    d.out = out.data_ptr()
    lib = L.lib()
    with torch.cuda.device(x.device):
        ws_bytes = lib.b200phi_reduce_workspace(C.byref(d))
        ws = None
        if ws_bytes:
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)  # dev_ctx.Alloc stand-in
            d.workspace = ws.data_ptr()
            d.workspace_bytes = ws_bytes
        L.check(lib.b200phi_reduce(C.byref(d), _stream_ptr(x)))
    return out


def _sum_out_dtype(x_dtype, dtype):
    """SumRawInferMeta, paddle/phi/infermeta/unary.cc:4698-4707."""
    if dtype is not None:
        return dtype
    if x_dtype in (torch.bool, torch.int32, torch.int64):
        return torch.int64
    return x_dtype


def sum(x, axis=None, dtype=None, keepdim=False, out=None):  # noqa: A001 - paddle name
    return _reduce(L.SUM, x, axis, keepdim, _sum_out_dtype(x.dtype, _to_torch_dtype(dtype)), out)


def mean(x, axis=None, keepdim=False):
    return _reduce(L.MEAN, x, axis, keepdim, x.dtype)


def max(x, axis=None, keepdim=False, out=None, dtype=None):  # noqa: A001
    """`dtype`: the C ABI's out_dtype (phi::Reduce casts x to it first, gpu/reduce.h:48-69)."""
    return _reduce(L.MAX, x, axis, keepdim, _to_torch_dtype(dtype) or x.dtype, out)


def min(x, axis=None, keepdim=False, out=None, dtype=None):  # noqa: A001
    return _reduce(L.MIN, x, axis, keepdim, _to_torch_dtype(dtype) or x.dtype, out)


def amax(x, axis=None, keepdim=False):
    return _reduce(L.AMAX, x, axis, keepdim, x.dtype)


def amin(x, axis=None, keepdim=False):
    return _reduce(L.AMIN, x, axis, keepdim, x.dtype)


def prod(x, axis=None, keepdim=False, dtype=None):
    dtype = _to_torch_dtype(dtype)
    if dtype is not None and dtype != x.dtype:
        # math.py:4779+: prod casts x first when dtype is given
        x = cast(x, dtype)
    return _reduce(L.PROD, x, axis, keepdim, x.dtype)


def any(x, axis=None, keepdim=False):  # noqa: A001
    return _reduce(L.ANY, x, axis, keepdim, torch.bool)


def all(x, axis=None, keepdim=False):  # noqa: A001
    return _reduce(L.ALL, x, axis, keepdim, torch.bool)


def cast(x, dtype):
    """phi::Cast for the dtypes the reduce engine reads: a reduce with
    reduce_num == 1 is the reference's own elementwise shortcut
    (reduce_function.h:1087-1092)."""
    dtype = _to_torch_dtype(dtype)
    _require_cuda(x)
    x = x.contiguous()
    out = torch.empty(x.shape, dtype=dtype, device=x.device)
    if x.numel() == 0:
        return out
    d = L.ReduceDesc()
    d.op = L.SUM
    d.in_dtype = _phi_dtype(x.dtype)
    d.out_dtype = _phi_dtype(dtype)
    d.rank = 2
    d.dims[0] = x.numel()
    d.dims[1] = 1
    d.naxes = 1
    d.axes[0] = 1
    d.x = x.data_ptr()
    d.out = out.data_ptr()
    with torch.cuda.device(x.device):
        L.check(L.lib().b200phi_reduce(C.byref(d), _stream_ptr(x)))
    return out


# --------------------------------------------------------------------------- arg


def _arg(op, x, axis, keepdim, dtype, want_value=False):
    _require_cuda(x)
    x = x.contiguous()
    dtype = _to_torch_dtype(dtype)
    if dtype not in (torch.int32, torch.int64):
        # search.py:175+: "dtype" must be int32 or int64 (None is rejected too)
        raise ValueError("the value of 'dtype' in argmax/argmin must be int32 or int64")
    if x.numel() == 0:
        raise ValueError("argmin/argmax input numel must > 0")
    rank = x.dim()
    flatten = axis is None
    if flatten:
        ax = 0
        out_shape = [1] * rank if keepdim else []
    else:
        if not isinstance(axis, int):
            raise TypeError("The type of 'axis' must be int or None in argmax/argmin")
        lo, hi = (-rank, rank) if rank > 0 else (-1, 1)
        if not (lo <= axis < hi):
            raise ValueError(f"'axis' must be in [-{rank}, {rank}), but received {axis}")
        ax = axis + rank if axis < 0 else axis
        if rank == 0:
            out_shape = []
        elif keepdim:
            out_shape = [1 if i == ax else s for i, s in enumerate(x.shape)]
        else:
            out_shape = [s for i, s in enumerate(x.shape) if i != ax]
    # ArgMinMaxInferMeta (unary.cc:257-360): int32 needs the reduced extent <= INT_MAX
    extent = x.numel() if flatten else (x.shape[ax] if rank else 1)
    if dtype == torch.int32 and extent > 2**31 - 1:
        raise ValueError("argmax/argmin with int32 indices needs the reduced extent <= INT_MAX")
    out = torch.empty(out_shape, dtype=dtype, device=x.device)
    val = torch.empty(out_shape, dtype=x.dtype, device=x.device) if want_value else None
    d = L.ArgReduceDesc()
    d.op = op
    d.in_dtype = _phi_dtype(x.dtype)
    d.index_dtype = _phi_dtype(dtype)
    d.rank = rank
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.axis = ax
    d.flatten = 1 if flatten else 0
    d.x = x.data_ptr()
    d.out = out.data_ptr()
    d.max_out = val.data_ptr() if want_value else None
    with torch.cuda.device(x.device):
        ws_bytes = L.lib().b200phi_argreduce_workspace(C.byref(d))
        if ws_bytes:   # flatten, few long rows, a non-last axis: the reduced extent is split across CTAs
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
            d.workspace, d.workspace_bytes = ws.data_ptr(), ws_bytes
        L.check(L.lib().b200phi_argreduce(C.byref(d), _stream_ptr(x)))
    return (out, val) if want_value else out


def argmax(x, axis=None, keepdim=False, dtype="int64"):
    return _arg(L.ARGMAX, x, axis, keepdim, dtype)


def argmin(x, axis=None, keepdim=False, dtype="int64"):
    return _arg(L.ARGMIN, x, axis, keepdim, dtype)


def argmax_with_value(x, axis=None, keepdim=False, dtype="int64"):
    """One pass over x for both paddle.argmax and paddle.max (config C3)."""
    return _arg(L.ARGMAX, x, axis, keepdim, dtype, want_value=True)


def argmin_with_value(x, axis=None, keepdim=False, dtype="int64"):
    return _arg(L.ARGMIN, x, axis, keepdim, dtype, want_value=True)


# --------------------------------------------------------------------- row stats


def row_mean_var(x, unbiased=True, out_dtype=None):
    """mean and variance over the last axis in a single read of x (config C4)."""
    _require_cuda(x)
    x = x.contiguous()
    if x.dim() < 1:
        raise ValueError("row_mean_var needs rank >= 1")
    cols = x.shape[-1]
    rows = x.numel() // cols if cols else 0
    out_dtype = _to_torch_dtype(out_dtype) or x.dtype
    m = torch.empty(x.shape[:-1], dtype=out_dtype, device=x.device)
    v = torch.empty(x.shape[:-1], dtype=out_dtype, device=x.device)
    d = L.RowStatsDesc()
    d.in_dtype = _phi_dtype(x.dtype)
    d.out_dtype = _phi_dtype(out_dtype)
    d.rows = rows
    d.cols = cols
    d.unbiased = 1 if unbiased else 0
    d.x = x.data_ptr()
    d.mean = m.data_ptr()
    d.var = v.data_ptr()
    with torch.cuda.device(x.device):
        L.check(L.lib().b200phi_rowstats(C.byref(d), _stream_ptr(x)))
    return m, v


def var(x, axis=None, unbiased=True, keepdim=False):
    """paddle.var (stat.py:141-198).  The last-axis case is one fused kernel; any
    other axis set runs the reference's own composite (mean -> subtract -> square
    -> sum -> divide) on the b200phi reduce and broadcast kernels."""
    _require_cuda(x)
    rank = x.dim()
    axes, reduce_all = _norm_axes(axis, rank)
    if x.dtype in (torch.float32, torch.float64, torch.float16, torch.bfloat16) and rank >= 1 \
            and not reduce_all and axes == [rank - 1]:
        _, v = row_mean_var(x, unbiased=unbiased)
        return v.unsqueeze(-1) if keepdim else v
    u = mean(x, axis, keepdim=True)
    diff = subtract(x, u)
    sq = multiply(diff, diff)
    ssum = sum(sq, axis, keepdim=keepdim)
    n = 1
    for a in (range(rank) if reduce_all else axes):
        n *= x.shape[a]
    if unbiased:
        n = n - 1 if n > 1 else 1  # stat.py:193-196: where(n > 1, n - 1, 1)
    return scale(ssum, 1.0 / n)


# ------------------------------------------------------------------- elementwise


def broadcast_shape(xs, ys, axis=-1):
    """Output dims of a binary op: paddle/phi/infermeta/binary.cc:1665-1740 with
    funcs::GetBroadcastDimsArrays (common_shape.h:35-92)."""
    xs, ys = list(xs), list(ys)
    if len(xs) == len(ys) and axis not in (-1, 0):
        raise ValueError("axis should be -1 or 0 while the dimension of tensor X == tensor Y")
    max_dim = len(xs) if len(xs) >= len(ys) else len(ys)
    if axis < -1 - max_dim or axis > max_dim:
        raise ValueError(f"The axis range must be [{-max_dim - 1}, {max_dim}], but axis is {axis}.")
    ax = (len(xs) - len(ys) if len(xs) >= len(ys) else len(ys) - len(xs)) + axis + 1 if axis < 0 else axis
    if len(xs) >= len(ys):
        big, small, small_is_y = xs, ys, True
    else:
        big, small, small_is_y = ys, xs, False
    if ax + len(small) > max_dim:
        raise ValueError("Broadcast dimension mismatch: axis places the smaller operand out of range")
    padded = [1] * ax + small + [1] * (max_dim - ax - len(small))
    out = []
    for a, b in zip(big, padded):
        if not (a == b or a <= 1 or b <= 1):
            x_, y_ = (big, padded) if small_is_y else (padded, big)
            raise ValueError(
                f"Broadcast dimension mismatch. Operands could not be broadcast together with "
                f"the shape of X = {x_} and the shape of Y = {y_}.")
        out.append(0 if (a == 0 or b == 0) else (a if a > b else b))
    return out


def _elementwise(binary, ins, act=L.ACT_NONE, compound=0, scale_v=0.0, axis=-1,
                 want_intermediate=False, out_shape=None):
    _require_cuda(*ins)
    ins = [t.contiguous() for t in ins]
    dt = ins[0].dtype
    for t in ins[1:]:
        if t.dtype != dt:
            raise TypeError("b200phi elementwise: all inputs must share one dtype "
                            "(type promotion happens above the kernel in Paddle)")
    if out_shape is not None:
        out_shape = list(out_shape)  # a unary op broadcast into a larger shape (reduce_grad.h:30-42)
    elif len(ins) == 1:
        out_shape = list(ins[0].shape)
    else:
        out_shape = broadcast_shape(ins[0].shape, ins[1].shape, axis)
        for t in ins[2:]:
            out_shape = broadcast_shape(out_shape, t.shape, -1)
    out = torch.empty(out_shape, dtype=dt, device=ins[0].device)
    inter = torch.empty(out_shape if compound == 0 else list(ins[1].shape), dtype=dt,
                        device=ins[0].device) if want_intermediate else None
    d = L.ElementwiseDesc()
    d.binary = binary
    d.act = act
    d.compound = compound
    d.scale = scale_v
    d.dtype = _phi_dtype(dt)
    d.arity = len(ins)
    d.axis = axis
    for i, t in enumerate(ins):
        d.in_[i].rank = t.dim()
        for j, s in enumerate(t.shape):
            d.in_[i].dims[j] = s
        d.in_[i].data = t.data_ptr()
    d.out_rank = len(out_shape)
    for j, s in enumerate(out_shape):
        d.out_dims[j] = s
    d.out = out.data_ptr()
    d.intermediate_out = inter.data_ptr() if want_intermediate else None
    if out.numel() > 0:
        with torch.cuda.device(out.device):
            L.check(L.lib().b200phi_elementwise(C.byref(d), _stream_ptr(out)))
    return (out, inter) if want_intermediate else out


def add(x, y, axis=-1, out=None):
    """paddle.add.  A float32 x with a bfloat16 / float16 y takes the mixed-precision route of
    phi::AddKernel (kps/elementwise_kernel.cu:82-99): out = x + widen(y) in float32, one launch,
    no cast pass; `out=x` accumulates in place (the master-gradient update)."""
    if (isinstance(x, torch.Tensor) and isinstance(y, torch.Tensor) and x.dtype == torch.float32
            and y.dtype in (torch.bfloat16, torch.float16)):
        return _mixed_add(x, y, out)
    if out is not None:
        raise ValueError("out= is only supported on the mixed-precision route")
    return _elementwise(L.ADD, [x, y], axis=axis)


def _mixed_add(x, y, out=None):
    _require_cuda(x, y)
    x, y = x.contiguous(), y.contiguous()
    out_shape = broadcast_shape(x.shape, y.shape, -1)
    if out is None:
        out = torch.empty(out_shape, dtype=torch.float32, device=x.device)
    else:
        _require_cuda(out)
        if list(out.shape) != out_shape or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous float32 tensor of shape {out_shape}")
        if out.data_ptr() == x.data_ptr() and list(x.shape) != out_shape:
            raise ValueError("in-place accumulation needs x to have the output shape")
    d = L.MixedAddDesc()
    d.y_dtype = _phi_dtype(y.dtype)
    for view, t in ((d.x, x), (d.y, y)):
        view.rank = t.dim()
        for j, s_ in enumerate(t.shape):
            view.dims[j] = s_
        view.data = t.data_ptr()
    d.out_rank = len(out_shape)
    for j, s_ in enumerate(out_shape):
        d.out_dims[j] = s_
    d.out = out.data_ptr()
    if out.numel() > 0:
        with torch.cuda.device(out.device):
            L.check(L.lib().b200phi_mixed_add(C.byref(d), _stream_ptr(out)))
    return out


def subtract(x, y, axis=-1):
    return _elementwise(L.SUBTRACT, [x, y], axis=axis)


def multiply(x, y, axis=-1):
    return _elementwise(L.MULTIPLY, [x, y], axis=axis)


def divide(x, y, axis=-1):
    return _elementwise(L.DIVIDE, [x, y], axis=axis)


def maximum(x, y, axis=-1):
    return _elementwise(L.MAXIMUM, [x, y], axis=axis)


def minimum(x, y, axis=-1):
    return _elementwise(L.MINIMUM, [x, y], axis=axis)


def add_n3(a, b, c):
    """Ternary broadcast add, the shape of test_ternary_broadcast.cu:96-120."""
    return _elementwise(L.ADD, [a, b, c])


def relu(x):
    return _elementwise(L.ADD, [x], act=L.ACT_RELU)


def scale(x, s):
    return _elementwise(L.ADD, [x], act=L.ACT_SCALE, scale_v=float(s))


_BINARY_NAMES = {"elementwise_add": L.ADD, "elementwise_sub": L.SUBTRACT,
                 "elementwise_mul": L.MULTIPLY}
_UNARY_NAMES = {"relu": L.ACT_RELU, "scale": L.ACT_SCALE, "gelu": L.ACT_GELU,
                "sigmoid": L.ACT_SIGMOID, "tanh": L.ACT_TANH}


def fused_elemwise_activation(x, y, functor_list, axis=-1, scale=0.0,  # noqa: A002
                              save_intermediate_out=False):
    """fused_elemwise_activation / fused_elemwise_add_activation
    (fused_ops.yaml:349-357; RunFunctors, fused_elemwise_activation_functor.h:388-458).
    functor_list = [outer, inner]:
      ["relu", "elementwise_add"]  -> out = relu(x + y),   intermediate = x + y
      ["elementwise_add", "relu"]  -> out = x + relu(y),   intermediate = relu(y)
    """
    if len(functor_list) != 2:
        raise ValueError("functor_list must hold exactly two functor names")
    f0, f1 = functor_list
    if f0 in _UNARY_NAMES and f1 in _BINARY_NAMES:
        compound, act, binary = 0, _UNARY_NAMES[f0], _BINARY_NAMES[f1]
    elif f0 in _BINARY_NAMES and f1 in _UNARY_NAMES:
        compound, act, binary = 1, _UNARY_NAMES[f1], _BINARY_NAMES[f0]
    else:
        raise ValueError(f"{functor_list} is not a (unary, binary) or (binary, unary) pair")
    return _elementwise(binary, [x, y], act=act, compound=compound, scale_v=float(scale),
                        axis=axis, want_intermediate=save_intermediate_out)


def fused_elemwise_add_activation(x, y, functor_list, axis=-1, scale=0.0,  # noqa: A002
                                  save_intermediate_out=False):
    if "elementwise_add" not in functor_list:
        raise ValueError("fused_elemwise_add_activation needs elementwise_add in functor_list")
    return fused_elemwise_activation(x, y, functor_list, axis, scale, save_intermediate_out)


# ------------------------------------------------------------------- functor plug-ins (map)


def _map(name, ins, p0=0.0, p1=0.0, axis=-1, out_shape=None):
    """One b200phi_map launch: compare / logical / bitwise / backward / activation functors on
    the broadcast engine (kps/compare_kernel.cu, kps/logical_kernel.cu, kps/bitwise_kernel.cu,
    gpu/activation_kernel.cu, gpu/elementwise_grad.h, gpu/reduce_kernel.cu:74-119)."""
    op = L.MAP_OPS[name]
    if len(ins) != L.map_arity(op):
        raise ValueError(f"{name} takes {L.map_arity(op)} operand(s)")
    _require_cuda(*ins)
    ins = [t.contiguous() for t in ins]
    dt = ins[0].dtype
    for t in ins[1:]:
        if t.dtype != dt:
            raise TypeError("b200phi map: all operands must share one dtype")
    if out_shape is not None:
        out_shape = list(out_shape)
    elif len(ins) == 1:
        out_shape = list(ins[0].shape)
    else:
        out_shape = broadcast_shape(ins[0].shape, ins[1].shape, axis)
        for t in ins[2:]:
            out_shape = broadcast_shape(out_shape, t.shape, -1)
    out = torch.empty(out_shape, dtype=torch.bool if L.map_out_is_bool(op) else dt,
                      device=ins[0].device)
    d = L.MapDesc()
    d.op = op
    d.in_dtype = _phi_dtype(dt)
    d.out_dtype = L.UNDEFINED
    d.arity = len(ins)
    d.axis = axis
    d.p0, d.p1 = float(p0), float(p1)
    for i, t in enumerate(ins):
        d.in_[i].rank = t.dim()
        for j, s in enumerate(t.shape):
            d.in_[i].dims[j] = s
        d.in_[i].data = t.data_ptr()
    d.out_rank = len(out_shape)
    for j, s in enumerate(out_shape):
        d.out_dims[j] = s
    d.out = out.data_ptr()
    if out.numel() > 0:
        with torch.cuda.device(out.device):
            L.check(L.lib().b200phi_map(C.byref(d), _stream_ptr(out)))
    return out


def _binary_map(name):
    def f(x, y, axis=-1):
        return _map(name, [x, y], axis=axis)
    f.__name__ = name
    f.__doc__ = f"paddle.{name}(x, y) through b200phi_map."
    return f


def _unary_map(name):
    def f(x):
        return _map(name, [x])
    f.__name__ = name
    f.__doc__ = f"paddle.{name}(x) through b200phi_map."
    return f


equal = _binary_map("equal")
not_equal = _binary_map("not_equal")
less_than = _binary_map("less_than")
less_equal = _binary_map("less_equal")
greater_than = _binary_map("greater_than")
greater_equal = _binary_map("greater_equal")
logical_and = _binary_map("logical_and")
logical_or = _binary_map("logical_or")
logical_xor = _binary_map("logical_xor")
logical_not = _unary_map("logical_not")
bitwise_and = _binary_map("bitwise_and")
bitwise_or = _binary_map("bitwise_or")
bitwise_xor = _binary_map("bitwise_xor")
bitwise_not = _unary_map("bitwise_not")
# the other arithmetic binaries of kps/elementwise_kernel.cu:183-293 (python/paddle/tensor/math.py:
# remainder / mod, floor_divide, pow(x, Tensor), copysign, fmax, fmin, heaviside)
remainder = _binary_map("remainder")
floor_divide = _binary_map("floor_divide")
elementwise_pow = _binary_map("elementwise_pow")
copysign = _binary_map("copysign")
fmax = _binary_map("fmax")
fmin = _binary_map("fmin")
heaviside = _binary_map("heaviside")


def bitwise_left_shift(x, y, is_arithmetic=True):
    """paddle.bitwise_left_shift (BitwiseLeftShiftKernel, kps/bitwise_kernel.cu)."""
    return _map("left_shift_arithmetic" if is_arithmetic else "left_shift_logic", [x, y])


def bitwise_right_shift(x, y, is_arithmetic=True):
    """paddle.bitwise_right_shift: sign-filling when is_arithmetic, zero-filling otherwise."""
    return _map("right_shift_arithmetic" if is_arithmetic else "right_shift_logic", [x, y])

exp = _unary_map("exp")
log = _unary_map("log")
sqrt = _unary_map("sqrt")
rsqrt = _unary_map("rsqrt")
square = _unary_map("square")
reciprocal = _unary_map("reciprocal")
sin = _unary_map("sin")
cos = _unary_map("cos")
tanh = _unary_map("tanh")
sigmoid = _unary_map("sigmoid")
silu = _unary_map("silu")
floor = _unary_map("floor")
ceil = _unary_map("ceil")
expm1 = _unary_map("expm1")
log1p = _unary_map("log1p")
log2 = _unary_map("log2")
log10 = _unary_map("log10")
softsign = _unary_map("softsign")
log_sigmoid = _unary_map("log_sigmoid")
abs = _unary_map("abs")  # noqa: A001 - paddle name


def leaky_relu(x, negative_slope=0.01):
    return _map("leaky_relu", [x], p0=negative_slope)


def elu(x, alpha=1.0):
    return _map("elu", [x], p0=alpha)


def softplus(x, beta=1.0, threshold=20.0):
    return _map("softplus", [x], p0=beta, p1=threshold)


def relu6(x, threshold=6.0):
    return _map("relu6", [x], p0=threshold)


def hardsigmoid(x, slope=0.1666667, offset=0.5):
    return _map("hard_sigmoid", [x], p0=slope, p1=offset)


def swish(x, beta=1.0):
    return _map("swish", [x], p0=beta)


def mish(x, threshold=20.0):
    return _map("mish", [x], p0=threshold)


# ------------------------------------------------------------------- reduce-side plug-ins
# The remaining clients of funcs::ReduceKernel named in SURVEY.md 8(f) row 2:
# gpu/{squared_l2_norm, p_norm, mean_all, logsumexp, dist, reduce_as}_kernel.cu and
# gpu/broadcast_tensors_kernel.cu.


def squared_l2_norm(x):
    """sum(x*x) over everything, shape [1] (SquaredL2NormKernel, gpu/squared_l2_norm_kernel.cu:24-34)."""
    return _reduce(L.SUM_SQUARE, x, None, False, x.dtype).reshape([1])


def p_norm(x, porder=2.0, axis=-1, epsilon=1e-12, keepdim=False, asvector=False):
    """PNormKernel (gpu/p_norm_kernel.cu:54-96): porder 0 counts non-zeros, +-inf is max/min |x|,
    otherwise (sum |x|^p)^(1/p) along `axis` (or over everything with asvector).  porder 1 and 2
    are single launches; any other finite p goes |x|^p -> sum -> ^(1/p), the same three
    functors the reference applies.  `epsilon` only matters to the backward kernel."""
    ax = None if asvector else [axis]
    porder = float(porder)
    if porder == 0:
        return _reduce(L.COUNT_NONZERO, x, ax, keepdim, x.dtype)
    if porder == float("inf"):
        return _reduce(L.MAX_ABS, x, ax, keepdim, x.dtype)
    if porder == float("-inf"):
        return _reduce(L.MIN_ABS, x, ax, keepdim, x.dtype)
    if porder == 1.0:
        return _reduce(L.SUM_ABS, x, ax, keepdim, x.dtype)
    if porder == 2.0:
        return _reduce(L.L2_NORM, x, ax, keepdim, x.dtype)
    s = _reduce(L.SUM, _map("pow_abs", [x], p0=porder), ax, keepdim, x.dtype)
    return _map("pow_abs", [s], p0=1.0 / porder)


def mean_all(x):
    """MeanAllKernel (gpu/mean_all_kernel.cu:27-53): mean over every element, 0-D result."""
    return _reduce(L.MEAN, x, None, False, x.dtype)


def reduce_as(x, target):
    """ReduceAsKernel (gpu/reduce_as_kernel.cu:24-36): sum x down to target's shape over the dims
    funcs::GetReduceDims finds (leading extra dims, and dims where target has 1)."""
    xs, ts = list(x.shape), list(target.shape)
    off = len(xs) - len(ts)
    if off < 0:
        raise ValueError("reduce_as: target has more dims than x")
    dims = list(range(off)) + [i + off for i, t in enumerate(ts) if t == 1 and xs[i + off] != 1]
    for i, t in enumerate(ts):
        if t != 1 and t != xs[i + off]:
            raise ValueError(f"reduce_as: target dim {i} = {t} does not divide x's shape {xs}")
    if not dims:
        return cast(x, x.dtype)
    return _reduce(L.SUM, x, dims, False, x.dtype).reshape(ts)


def dist(x, y, p=2.0):
    """DistKernel (gpu/dist_kernel.cu): the p-norm of (x - y) over every element, 0-D result."""
    return p_norm(subtract(x, y), porder=p, asvector=True).reshape([])


def broadcast_tensors(inputs):
    """BroadcastTensorsKernel (gpu/broadcast_tensors_kernel.cu): every input expanded to the common
    broadcast shape, each through one unary launch of the broadcast engine."""
    shape = list(inputs[0].shape)
    for t in inputs[1:]:
        shape = broadcast_shape(shape, list(t.shape), -1)
    return [_elementwise(L.ADD, [t], out_shape=shape) for t in inputs]


def equal_all(x, y):
    """EqualAllKernel (kps/compare_kernel.cu:78-103): false when the shapes differ, else
    all(x == y) as a 0-D bool."""
    if list(x.shape) != list(y.shape):
        return torch.zeros([], dtype=torch.bool, device=x.device)
    if x.numel() == 0:
        return torch.ones([], dtype=torch.bool, device=x.device)
    return _reduce(L.ALL, _map("equal", [x, y]), None, False, torch.bool)


def logsumexp(x, axis=None, keepdim=False):
    """paddle.logsumexp (LogsumexpKernel, gpu/logsumexp_kernel.cu:81-165).  Trailing (or all) axes:
    one launch of the row kernel, one read of x.  Other axis sets: max -> exp(x - max) -> sum ->
    log + max, the reference's own fallback (:49-78) on the two engines."""
    _require_cuda(x)
    x = x.contiguous()
    rank = x.dim()
    axes, reduce_all = _norm_axes(axis, rank)
    out_shape = _reduce_out_shape(list(x.shape), axes, reduce_all, keepdim)
    if x.numel() == 0:
        raise L.B200PhiError(1, "The dims of Input(X) should be greater than 0.")
    red = sorted(range(rank)) if reduce_all else sorted(axes)
    trailing = reduce_all or red == list(range(rank - len(red), rank))
    if trailing:
        cols = 1
        for a in red:
            cols *= x.shape[a]
        rows = x.numel() // cols
        out = torch.empty(out_shape, dtype=x.dtype, device=x.device)
        d = L.LogsumexpDesc()
        d.dtype = _phi_dtype(x.dtype)
        d.rows, d.cols = rows, cols
        d.x, d.out = x.data_ptr(), out.data_ptr()
        with torch.cuda.device(x.device):
            L.check(L.lib().b200phi_logsumexp(C.byref(d), _stream_ptr(x)))
        return out
    m = _reduce(L.MAX, x, red, True, x.dtype)
    s = _reduce(L.SUM, _map("sub_exp", [x, m]), red, True, x.dtype)
    return add(_map("log", [s]), m).reshape(out_shape)
