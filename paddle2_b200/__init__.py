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
from ._lib import B200PhiError  # noqa: F401
