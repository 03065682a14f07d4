"""TEST INFRASTRUCTURE -- numpy front-end of the two CPU checkers.

  port()  oracle/libphi_oracle.so   plain-C restatement (oracle/phi_oracle.c)
  ref()   oracle/_ref/libphi_ref.so the reference's own functors + vendored Eigen
                                    (oracle/ref_src/*.cc compiled against /root/reference)

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference)
import this module.  Nothing in paddle2_b200/ does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_PATH = os.path.join(_HERE, "libphi_oracle.so")
REF_PATH = os.path.join(_HERE, "_ref", "libphi_ref.so")

BOOL, UINT8, INT8, INT16, INT32, INT64, F32, F64 = 1, 2, 3, 5, 7, 9, 10, 11
_NP2DT = {np.dtype(np.bool_): BOOL, np.dtype(np.uint8): UINT8, np.dtype(np.int8): INT8,
          np.dtype(np.int16): INT16, np.dtype(np.int32): INT32, np.dtype(np.int64): INT64,
          np.dtype(np.float32): F32, np.dtype(np.float64): F64}
OPS = {"sum": 0, "mean": 1, "max": 2, "min": 3, "prod": 4, "any": 5, "all": 6, "amax": 7, "amin": 8}
BINARY = {"add": 0, "subtract": 1, "multiply": 2, "divide": 3, "maximum": 4, "minimum": 5}
UNARY = {None: 0, "none": 0, "relu": 1, "scale": 2, "gelu": 3, "sigmoid": 4, "tanh": 5}

_port = None
_ref = None


def _c(x, dtype=None):
    """C-contiguous array that keeps 0-D arrays 0-D (np.ascontiguousarray would make them 1-D)."""
    return np.require(np.asarray(x, dtype=dtype), requirements=["C"])


def build():
    """Compile the port (and, where /root/reference exists, the reference build)."""
    import subprocess
    subprocess.run(["make", "-s", "-j8", "-C", _HERE, "all"], check=True)


def port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_PATH):
            build()
        _port = C.CDLL(PORT_PATH)
    return _port


def have_ref():
    return os.path.exists(REF_PATH)


def ref():
    global _ref
    if _ref is None:
        _ref = C.CDLL(REF_PATH)
    return _ref


def bf16_bits_to_f32(bits):
    """uint16 bf16 bit patterns -> float32 (exact)."""
    bits = _c(bits, dtype=np.uint16)
    return (bits.astype(np.uint32) << 16).view(np.float32)


def f32_to_bf16_bits(x):
    """float32 -> bf16 bit patterns, round-to-nearest-even."""
    u = _c(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = ((u >> 16) & 1) + 0x7FFF
    return ((u + r) >> 16).astype(np.uint16)


def _axes_list(axis, rank):
    if axis is None:
        return [], True
    if isinstance(axis, (int, np.integer)):
        axis = [int(axis)]
    axis = [int(a) for a in axis]
    if len(axis) == 0:
        return [], True
    norm = sorted({a + rank if a < 0 else a for a in axis})
    return norm, len(norm) == rank or rank == 0


def _out_shape(shape, axes, reduce_all, keepdim):
    if reduce_all:
        return [1] * len(shape) if keepdim else []
    if keepdim:
        return [1 if i in axes else s for i, s in enumerate(shape)]
    return [s for i, s in enumerate(shape) if i not in axes]


def sum_out_dtype(x_dtype, dtype=None):
    """SumRawInferMeta (infermeta/unary.cc:4698-4707)."""
    if dtype is not None:
        return np.dtype(dtype)
    if np.dtype(x_dtype) in (np.dtype(np.bool_), np.dtype(np.int32), np.dtype(np.int64)):
        return np.dtype(np.int64)
    return np.dtype(x_dtype)


def reduce(op, x, axis=None, keepdim=False, out_dtype=None, impl="port"):
    """Reduce like phi::Reduce<CPUContext> does. impl: 'port' or 'ref'."""
    x = _c(x)
    rank = x.ndim
    axes, reduce_all = _axes_list(axis, rank)
    logical = op in ("any", "all")
    if out_dtype is None:
        out_dtype = np.dtype(np.bool_) if logical else x.dtype
    out_dtype = np.dtype(out_dtype)
    out = np.empty(_out_shape(list(x.shape), axes, reduce_all, keepdim), dtype=out_dtype)
    dims = (C.c_int64 * 9)(*x.shape)
    if impl == "port":
        ax = (C.c_int32 * 9)(*axes)
        rc = port().phi_oracle_reduce(OPS[op], _NP2DT[x.dtype], _NP2DT[out_dtype], rank, dims,
                                      len(axes), ax, int(reduce_all),
                                      x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise ValueError(f"phi_oracle_reduce status {rc}")
        return out
    # reference build: cast first (phi::Cast), then the typed functor
    xin = x.astype(out_dtype) if x.dtype != out_dtype else x
    if logical:
        xin = (x != 0)
    xin = _c(xin)
    names = {np.dtype(np.float32): "float", np.dtype(np.float64): "double",
             np.dtype(np.int32): "int32", np.dtype(np.int64): "int64", np.dtype(np.bool_): "bool"}
    opname = {"amax": "max", "amin": "min"}.get(op, op)
    fn = getattr(ref(), f"phi_ref_{opname}_{names[xin.dtype]}")
    ax = (C.c_int64 * 9)(*axes)
    rc = fn(xin.ctypes.data_as(C.c_void_p), rank, dims, len(axes), ax, int(reduce_all),
            out.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise ValueError(f"phi_ref reduce status {rc}")
    return out


def argreduce(op, x, axis=None, keepdim=False, dtype=np.int64, impl="port"):
    """argmax/argmin like the CPU kernel (Eigen argmax(axis).cast<Tout>())."""
    x = _c(x)
    rank = x.ndim
    flatten = axis is None
    ax = 0 if flatten else (axis + rank if axis < 0 else axis)
    if flatten or rank == 0:
        shape = [1] * rank if keepdim else []
    elif keepdim:
        shape = [1 if i == ax else s for i, s in enumerate(x.shape)]
    else:
        shape = [s for i, s in enumerate(x.shape) if i != ax]
    out = np.empty(shape, dtype=np.int64)
    dims = (C.c_int64 * 9)(*x.shape)
    is_max = op == "argmax"
    if impl == "port":
        rc = port().phi_oracle_argreduce(0 if is_max else 1, _NP2DT[x.dtype], rank, dims,
                                         C.c_int64(ax), int(flatten),
                                         x.ctypes.data_as(C.c_void_p),
                                         out.ctypes.data_as(C.POINTER(C.c_int64)))
    else:
        names = {np.dtype(np.float32): "float", np.dtype(np.float64): "double",
                 np.dtype(np.int32): "int32", np.dtype(np.int64): "int64"}
        fn = getattr(ref(), f"phi_ref_arg_{names[x.dtype]}")
        rc = fn(x.ctypes.data_as(C.c_void_p), rank, dims, C.c_int64(ax), int(flatten), int(is_max),
                out.ctypes.data_as(C.POINTER(C.c_int64)))
    if rc != 0:
        raise ValueError(f"oracle argreduce status {rc}")
    return out.astype(dtype)


def _padded_dims(shape, out_rank, offset):
    d = [1] * out_rank
    for k, s in enumerate(shape):
        d[offset + k] = s
    return d


def broadcast_out_shape(xs, ys, axis=-1):
    """GetBroadcastDimsArrays (common_shape.h:35-92) + ElementwiseInferMeta (binary.cc:1665-1740)."""
    xs, ys = list(xs), list(ys)
    R = max(len(xs), len(ys))
    ax = abs(len(xs) - len(ys)) if axis < 0 else axis
    if len(xs) >= len(ys):
        xp, yp = _padded_dims(xs, R, 0), _padded_dims(ys, R, ax if len(ys) < len(xs) else 0)
    else:
        xp, yp = _padded_dims(xs, R, ax), _padded_dims(ys, R, 0)
    out = []
    for a, b in zip(xp, yp):
        if not (a == b or a <= 1 or b <= 1):
            raise ValueError("Broadcast dimension mismatch.")
        out.append(0 if (a == 0 or b == 0) else max(a, b))
    return out, xp, yp


def elementwise(binary, ins, act=None, compound=0, scale=0.0, axis=-1, want_intermediate=False,
                impl="port"):
    """out = act(binary(x, y[, z])) (compound 0) or binary(x, act(y)) (compound 1)."""
    ins = [_c(t) for t in ins]
    dt = ins[0].dtype
    if len(ins) == 1:
        out_shape, pads = list(ins[0].shape), [list(ins[0].shape)]
    else:
        out_shape, xp, yp = broadcast_out_shape(ins[0].shape, ins[1].shape, axis)
        pads = [xp, yp]
        for t in ins[2:]:
            prev = out_shape
            out_shape, _, zp = broadcast_out_shape(prev, t.shape, -1)
            pads = [_padded_dims(p, len(out_shape), len(out_shape) - len(p)) for p in pads] + [zp]
    R = len(out_shape)
    out = np.empty(out_shape, dtype=dt)
    inter = np.empty(out_shape, dtype=dt) if want_intermediate else None
    od = (C.c_int64 * 9)(*out_shape)
    ptrs = (C.c_void_p * 3)(*[t.ctypes.data for t in ins] + [None] * (3 - len(ins)))
    ip = inter.ctypes.data_as(C.c_void_p) if want_intermediate else None
    if impl == "port":
        flat = (C.c_int64 * 27)()
        for i, p in enumerate(pads):
            for k, s in enumerate(p):
                flat[i * 9 + k] = s
        rc = port().phi_oracle_elementwise(BINARY[binary], UNARY[act], int(compound),
                                           C.c_float(scale), _NP2DT[dt], len(ins), R, od, flat,
                                           ptrs, out.ctypes.data_as(C.c_void_p), ip)
    else:
        names = {np.dtype(np.float32): "float", np.dtype(np.float64): "double",
                 np.dtype(np.int32): "int32", np.dtype(np.int64): "int64"}
        rows = [(C.c_int64 * 9)(*p) for p in pads]
        rowptrs = (C.POINTER(C.c_int64) * 3)(*[C.cast(r, C.POINTER(C.c_int64)) for r in rows] +
                                             [None] * (3 - len(rows)))
        fn = getattr(ref(), f"phi_ref_elementwise_{names[dt]}")
        rc = fn(BINARY[binary], UNARY[act], int(compound), C.c_float(scale), len(ins), R, od,
                rowptrs, ptrs, out.ctypes.data_as(C.c_void_p), ip)
    if rc != 0:
        raise ValueError(f"oracle elementwise status {rc}")
    return (out, inter) if want_intermediate else out


MAP_OPS = {
    "equal": 0, "not_equal": 1, "less_than": 2, "less_equal": 3, "greater_than": 4,
    "greater_equal": 5,
    "logical_and": 16, "logical_or": 17, "logical_xor": 18, "logical_not": 19,
    "bitwise_and": 32, "bitwise_or": 33, "bitwise_xor": 34, "bitwise_not": 35,
    "left_shift_arithmetic": 36, "left_shift_logic": 37, "right_shift_arithmetic": 38,
    "right_shift_logic": 39,
    "equal_mul": 48, "max_grad_x": 49, "max_grad_y": 50, "min_grad_x": 51, "min_grad_y": 52,
    "div_grad_y": 53, "heaviside": 54, "equal_t": 56, "sub_exp": 57,
    "remainder": 58, "floor_divide": 59, "elementwise_pow": 60, "copysign": 61, "fmax": 62, "fmin": 63,
    "exp": 64, "log": 65, "sqrt": 66, "rsqrt": 67, "square": 68, "reciprocal": 69, "sin": 70,
    "cos": 71, "tanh": 72, "sigmoid": 73, "silu": 74, "leaky_relu": 75, "elu": 76,
    "softplus": 77, "floor": 78, "ceil": 79, "expm1": 80, "log1p": 81, "log2": 82, "log10": 83,
    "relu6": 84, "hard_sigmoid": 85, "swish": 86, "mish": 87, "softsign": 88,
    "log_sigmoid": 89, "abs": 90, "pow_abs": 91, "relu": 92,
}
_REF_NAMES = {np.dtype(np.float32): "float", np.dtype(np.float64): "double",
              np.dtype(np.int32): "int32", np.dtype(np.int64): "int64",
              np.dtype(np.int16): "int16", np.dtype(np.int8): "int8",
              np.dtype(np.uint8): "uint8", np.dtype(np.bool_): "bool"}


def map_arity(op):
    if op in (19, 35) or op >= 64:
        return 1
    return 3 if 48 <= op <= 53 else 2


def map(name, ins, p0=0.0, p1=0.0, axis=-1, impl="port"):  # noqa: A001
    """One functor plug-in of b200phi_map on numpy operands (compare / logical / bitwise /
    backward functors / activations).  impl="ref" runs the reference's own functors."""
    op = MAP_OPS[name]
    ins = [_c(t) for t in ins]
    if len(ins) != map_arity(op):
        raise ValueError(f"{name} takes {map_arity(op)} operand(s)")
    dt = ins[0].dtype
    if len(ins) == 1:
        out_shape, pads = list(ins[0].shape), [list(ins[0].shape)]
    else:
        out_shape, xp, yp = broadcast_out_shape(ins[0].shape, ins[1].shape, axis)
        pads = [xp, yp]
        for t in ins[2:]:
            prev = out_shape
            out_shape, _, zp = broadcast_out_shape(prev, t.shape, -1)
            pads = [_padded_dims(p, len(out_shape), len(out_shape) - len(p)) for p in pads] + [zp]
    R = len(out_shape)
    out = np.empty(out_shape, dtype=np.bool_ if op < 32 else dt)
    od = (C.c_int64 * 9)(*out_shape)
    ptrs = (C.c_void_p * 3)(*[t.ctypes.data for t in ins] + [None] * (3 - len(ins)))
    if impl == "port":
        flat = (C.c_int64 * 27)()
        for i, p in enumerate(pads):
            for k, s in enumerate(p):
                flat[i * 9 + k] = s
        rc = port().phi_oracle_map(op, _NP2DT[dt], C.c_float(p0), C.c_float(p1), len(ins), R, od,
                                   flat, ptrs, out.ctypes.data_as(C.c_void_p))
    elif op >= 64:
        fn = getattr(ref(), f"phi_ref_act_{_REF_NAMES[dt]}")
        rc = fn(op, C.c_float(p0), C.c_float(p1), ins[0].ctypes.data_as(C.c_void_p),
                out.ctypes.data_as(C.c_void_p), C.c_int64(ins[0].size))
    else:
        rows = [(C.c_int64 * 9)(*p) for p in pads]
        rowptrs = (C.POINTER(C.c_int64) * 3)(*[C.cast(r, C.POINTER(C.c_int64)) for r in rows] +
                                             [None] * (3 - len(rows)))
        fn = getattr(ref(), f"phi_ref_map_{_REF_NAMES[dt]}")
        rc = fn(op, R, od, rowptrs, ptrs, out.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise ValueError(f"oracle map status {rc} for {name} on {dt}")
    return out


def _pre_n_post(shape, axis, asvector):
    """GetDims of cpu/p_norm_kernel.cc:25-44."""
    if asvector or len(shape) == 0:
        return 1, int(np.prod(shape, dtype=np.int64)), 1
    axis = axis + len(shape) if axis < 0 else axis
    pre = int(np.prod(shape[:axis], dtype=np.int64))
    post = int(np.prod(shape[axis + 1:], dtype=np.int64))
    return pre, int(shape[axis]), post


def p_norm(x, porder=2.0, axis=-1, keepdim=False, asvector=False, impl="port"):
    """PNormKernel on the CPU (cpu/p_norm_kernel.cc:46-99) over the (pre, n, post) view."""
    x = _c(x)
    pre, n, post = _pre_n_post(list(x.shape), axis, asvector)
    if asvector or x.ndim == 0:
        out_shape = [1] * x.ndim if keepdim else []
    else:
        a = axis + x.ndim if axis < 0 else axis
        out_shape = [1 if i == a else s for i, s in enumerate(x.shape)] if keepdim else \
            [s for i, s in enumerate(x.shape) if i != a]
    if impl == "ref":
        out = np.empty((pre, post), dtype=x.dtype)
        fn = getattr(ref(), f"phi_ref_pnorm_{_REF_NAMES[x.dtype]}")
        rc = fn(x.ctypes.data_as(C.c_void_p), pre, n, post, C.c_float(porder),
                out.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise ValueError(f"oracle p_norm status {rc}")
        return out.reshape(out_shape)
    v = x.reshape(pre, n, post)
    if porder == 0:
        out = reduce("sum", (v != 0).astype(x.dtype), [1])
    elif porder == float("inf"):
        out = reduce("max", np.abs(v), [1])
    elif porder == float("-inf"):
        out = reduce("min", np.abs(v), [1])
    else:
        s = reduce("sum", map("pow_abs", [v], p0=porder), [1])
        out = map("pow_abs", [s], p0=1.0 / porder)
    return np.asarray(out).reshape(out_shape)


def squared_l2_norm(x, impl="port"):
    """input.square().sum() (funcs/squared_l2_norm.h:40-46): a full reduction of the squares."""
    x = _c(x)
    if impl == "ref":
        out = np.empty([1], dtype=x.dtype)
        fn = getattr(ref(), f"phi_ref_squared_l2_norm_{_REF_NAMES[x.dtype]}")
        rc = fn(x.ctypes.data_as(C.c_void_p), C.c_int64(x.size), out.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise ValueError(f"oracle squared_l2_norm status {rc}")
        return out
    return np.asarray(reduce("sum", (x * x).reshape(-1), None)).reshape([1])


def logsumexp(x, axis=None, keepdim=False, impl="port"):
    """LogsumexpFunctor (impl/logsumexp_kernel_impl.h:33-61): max + log(sum(exp(x - max)))."""
    x = _c(x)
    axes, reduce_all = _axes_list(axis, x.ndim)
    out_shape = _out_shape(list(x.shape), axes, reduce_all, keepdim)
    if impl == "ref":
        red = list(range(x.ndim)) if reduce_all else sorted(axes)
        if red != list(range(red[0], red[0] + len(red))) if red else False:
            raise ValueError("ref logsumexp: adjacent reduced axes only")
        pre = int(np.prod(x.shape[:red[0]], dtype=np.int64)) if red else 1
        n = int(np.prod([x.shape[a] for a in red], dtype=np.int64)) if red else 1
        post = x.size // max(pre * n, 1)
        out = np.empty((pre, post), dtype=x.dtype)
        fn = getattr(ref(), f"phi_ref_logsumexp_{_REF_NAMES[x.dtype]}")
        rc = fn(x.ctypes.data_as(C.c_void_p), pre, n, post, out.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise ValueError(f"oracle logsumexp status {rc}")
        return out.reshape(out_shape)
    ax = None if reduce_all else axes
    m = reduce("max", x, ax, keepdim=True)
    s = reduce("sum", map("sub_exp", [x, np.asarray(m)]), ax, keepdim=True)
    return (map("log", [np.asarray(s)]) + np.asarray(m)).astype(x.dtype).reshape(out_shape)


def mixed_add(x, y_bits, y_is_bf16=True, impl="port"):
    """fp32 x + widen(16-bit y): MultiPrecisionAddFunctor (elementwise_functor.h:45-50).  y is
    given as raw uint16 patterns (numpy has no bfloat16); same shape as x."""
    x = _c(x, dtype=np.float32)
    y_bits = _c(y_bits, dtype=np.uint16)
    if impl == "ref":
        out = np.empty_like(x)
        rc = ref().phi_ref_mixed_add(int(y_is_bf16), x.ctypes.data_as(C.c_void_p),
                                     y_bits.ctypes.data_as(C.c_void_p), C.c_int64(x.size),
                                     out.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise ValueError(f"oracle mixed_add status {rc}")
        return out
    yf = bf16_bits_to_f32(y_bits) if y_is_bf16 else y_bits.view(np.float16).astype(np.float32)
    return (x + yf.reshape(y_bits.shape)).astype(np.float32)


def rowstats_f32(x, unbiased=True):
    """(mean, var) over the last axis of a float32 array: paddle.var's composite in fp32."""
    x = _c(x, dtype=np.float32)
    cols = x.shape[-1]
    rows = x.size // cols
    mean = np.empty(x.shape[:-1], dtype=np.float32)
    var = np.empty(x.shape[:-1], dtype=np.float32)
    fn = port().phi_oracle_rowstats_f32
    rc = fn(C.c_int64(rows), C.c_int64(cols), int(unbiased), x.ctypes.data_as(C.c_void_p),
            mean.ctypes.data_as(C.c_void_p), var.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise ValueError(f"oracle rowstats status {rc}")
    return mean, var


def var_composite(x, axis=None, unbiased=True, keepdim=False, impl="port"):
    """paddle.var as the reference composes it (stat.py:184-197), every step through
    the oracle's reduce / elementwise, in x's own dtype (fp32 / fp64)."""
    x = _c(x)
    u = reduce("mean", x, axis, keepdim=True, impl=impl)
    d = elementwise("subtract", [x, u], impl=impl)
    sq = elementwise("multiply", [d, d], impl=impl)
    s = reduce("sum", sq, axis, keepdim=keepdim, impl=impl)
    axes, reduce_all = _axes_list(axis, x.ndim)
    n = x.size if reduce_all else int(np.prod([x.shape[a] for a in axes]))
    if unbiased:
        n = n - 1 if n > 1 else 1
    return (s / np.asarray(n, dtype=x.dtype)).astype(x.dtype)


def sum_axis0_f32_mt(x, threads):
    """bench.py only: axis-0 sum of a [rows, cols] fp32 array on `threads` host threads."""
    x = _c(x, dtype=np.float32)
    rows, cols = x.shape
    out = np.empty(cols, dtype=np.float32)
    rc = port().phi_oracle_sum_axis0_f32_mt(C.c_int64(rows), C.c_int64(cols),
                                            x.ctypes.data_as(C.c_void_p),
                                            out.ctypes.data_as(C.c_void_p), int(threads))
    if rc != 0:
        raise ValueError("oracle sum_axis0 status %d" % rc)
    return out
