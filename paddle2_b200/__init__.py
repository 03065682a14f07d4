"""paddle2_b200 -- B200-native phi reduce / broadcast hot path.

Only what the path needs lives here:
  csrc/      hand-written sm_100a CUDA kernels, the host planners (C++) and the C ABI
  _lib.py    ctypes binding of the C ABI (include/b200phi.h)
  ops.py     paddle.sum / max / argmax / var / add / fused_elemwise_add_activation look-alikes
  dist.py    sharded global reduce across the GPUs of one box (one process per GPU)
"""
from . import _lib  # noqa: F401
from .ops import (  # noqa: F401
    add, add_n3, all, amax, amin, any, argmax, argmax_with_value, argmin, argmin_with_value,
    broadcast_shape, cast, divide, fused_elemwise_activation, fused_elemwise_add_activation, max,
    maximum, mean, min, minimum, multiply, prod, relu, row_mean_var, scale, subtract, sum, var,
)
from .ops import (  # noqa: F401  -- functor plug-ins on the broadcast engine (b200phi_map)
    abs, bitwise_and, bitwise_left_shift, bitwise_not, bitwise_or, bitwise_right_shift, bitwise_xor, ceil, cos, elu, equal, exp, expm1, floor,
    greater_equal, greater_than, hardsigmoid, leaky_relu, less_equal, less_than, log, log10, log1p,
    log2, log_sigmoid, logical_and, logical_not, logical_or, logical_xor, mish, not_equal,
    reciprocal, relu6, rsqrt, sigmoid, silu, sin, softplus, softsign, sqrt, square, swish, tanh,
    copysign, elementwise_pow, floor_divide, fmax, fmin, heaviside, remainder,
)
from ._lib import B200PhiError  # noqa: F401
