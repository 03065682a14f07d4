"""ctypes binding of libb200phi.so (the C ABI in include/b200phi.h).

This is the *second* driver of the C ABI next to the phi adapter
(adapter/phi_b200_kernels.cu): tests and bench.py call the library the same way a
cgo/JNI/ctypes host would.  There is no CPU fallback anywhere in this package:
if the shared library is missing or a tensor is not on a CUDA device the call
raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200phi.so")

MAX_RANK = 9
MAX_ARITY = 3
NOALLOC_SCRATCH_BYTES = 4 << 20  # B200PHI_NOALLOC_SCRATCH_BYTES

# phi::DataType values (paddle/phi/common/data_type.h:42-85)
UNDEFINED, BOOL, UINT8, INT8, INT16, INT32, INT64 = 0, 1, 2, 3, 5, 7, 9
FLOAT32, FLOAT64, FLOAT16, BFLOAT16 = 10, 11, 15, 16
COMPLEX64, COMPLEX128, FLOAT8_E4M3FN, FLOAT8_E5M2 = 12, 13, 17, 18

SUM, MEAN, MAX, MIN, PROD, ANY, ALL, AMAX, AMIN = range(9)
SUM_ABS, SUM_SQUARE, L2_NORM, MAX_ABS, MIN_ABS, COUNT_NONZERO = range(9, 15)
ARGMAX, ARGMIN = 0, 1
ADD, SUBTRACT, MULTIPLY, DIVIDE, MAXIMUM, MINIMUM = range(6)
ACT_NONE, ACT_RELU, ACT_SCALE, ACT_GELU, ACT_SIGMOID, ACT_TANH = range(6)

# b200phi_map_op
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


def map_arity(op):
    """Operand count of a b200phi_map_op (include/b200phi.h)."""
    if op in (19, 35) or op >= 64:
        return 1
    return 3 if 48 <= op <= 53 else 2


def map_out_is_bool(op):
    return op < 32

STATUS_NAMES = {
    0: "OK",
    1: "INVALID_ARGUMENT",
    2: "UNSUPPORTED",
    3: "CUDA_ERROR",
    4: "WORKSPACE_TOO_SMALL",
    5: "NCCL_ERROR",
    6: "PEER_TIMEOUT",
}


class ReduceDesc(C.Structure):
    _fields_ = [
        ("op", C.c_int32),
        ("in_dtype", C.c_int32),
        ("out_dtype", C.c_int32),
        ("rank", C.c_int32),
        ("dims", C.c_int64 * MAX_RANK),
        ("naxes", C.c_int32),
        ("axes", C.c_int32 * MAX_RANK),
        ("reduce_all", C.c_int32),
        ("x", C.c_void_p),
        ("out", C.c_void_p),
        ("workspace", C.c_void_p),
        ("workspace_bytes", C.c_size_t),
    ]


class ArgReduceDesc(C.Structure):
    _fields_ = [
        ("op", C.c_int32),
        ("in_dtype", C.c_int32),
        ("index_dtype", C.c_int32),
        ("rank", C.c_int32),
        ("dims", C.c_int64 * MAX_RANK),
        ("axis", C.c_int64),
        ("flatten", C.c_int32),
        ("x", C.c_void_p),
        ("out", C.c_void_p),
        ("max_out", C.c_void_p),
        ("workspace", C.c_void_p),
        ("workspace_bytes", C.c_size_t),
    ]


class RowStatsDesc(C.Structure):
    _fields_ = [
        ("in_dtype", C.c_int32),
        ("out_dtype", C.c_int32),
        ("rows", C.c_int64),
        ("cols", C.c_int64),
        ("unbiased", C.c_int32),
        ("x", C.c_void_p),
        ("mean", C.c_void_p),
        ("var", C.c_void_p),
    ]


class TensorView(C.Structure):
    _fields_ = [
        ("rank", C.c_int32),
        ("dims", C.c_int64 * MAX_RANK),
        ("data", C.c_void_p),
    ]


class ElementwiseDesc(C.Structure):
    _fields_ = [
        ("binary", C.c_int32),
        ("act", C.c_int32),
        ("compound", C.c_int32),
        ("scale", C.c_float),
        ("dtype", C.c_int32),
        ("arity", C.c_int32),
        ("axis", C.c_int32),
        ("in_", TensorView * MAX_ARITY),
        ("out_rank", C.c_int32),
        ("out_dims", C.c_int64 * MAX_RANK),
        ("out", C.c_void_p),
        ("intermediate_out", C.c_void_p),
    ]


class LogsumexpDesc(C.Structure):
    _fields_ = [
        ("dtype", C.c_int32),
        ("rows", C.c_int64),
        ("cols", C.c_int64),
        ("x", C.c_void_p),
        ("out", C.c_void_p),
    ]


class MixedAddDesc(C.Structure):
    _fields_ = [
        ("y_dtype", C.c_int32),
        ("x", TensorView),
        ("y", TensorView),
        ("out_rank", C.c_int32),
        ("out_dims", C.c_int64 * MAX_RANK),
        ("out", C.c_void_p),
    ]


class MapDesc(C.Structure):
    _fields_ = [
        ("op", C.c_int32),
        ("in_dtype", C.c_int32),
        ("out_dtype", C.c_int32),
        ("arity", C.c_int32),
        ("axis", C.c_int32),
        ("p0", C.c_float),
        ("p1", C.c_float),
        ("in_", TensorView * MAX_ARITY),
        ("out_rank", C.c_int32),
        ("out_dims", C.c_int64 * MAX_RANK),
        ("out", C.c_void_p),
    ]


class P2PGroup(C.Structure):
    _fields_ = [
        ("nranks", C.c_int32),
        ("rank", C.c_int32),
        ("base", C.c_void_p * 8),
        ("max_out_numel", C.c_int64),
        ("max_tiles", C.c_int64),
    ]


class B200PhiError(RuntimeError):
    """A non-zero b200phi_status; mirrors the reference's EnforceNotMet."""

    def __init__(self, status, message):
        self.status = status
        super().__init__(f"b200phi {STATUS_NAMES.get(status, status)}: {message}")


# every symbol include/b200phi.h declares; tests check the library exports them all
EXPORTS = [
    "b200phi_reduce_workspace",
    "b200phi_reduce",
    "b200phi_reduce_plan_kind",
    "b200phi_reduce_noalloc",
    "b200phi_argreduce_workspace",
    "b200phi_argreduce",
    "b200phi_rowstats",
    "b200phi_elementwise",
    "b200phi_map",
    "b200phi_logsumexp",
    "b200phi_mixed_add",
    "b200phi_comm_unique_id",
    "b200phi_comm_init_rank",
    "b200phi_comm_destroy",
    "b200phi_sharded_reduce",
    "b200phi_sharded_reduce_workspace",
    "b200phi_sharded_argreduce_workspace",
    "b200phi_sharded_argreduce",
    "b200phi_p2p_buffer_bytes",
    "b200phi_p2p_alloc",
    "b200phi_p2p_open",
    "b200phi_p2p_close",
    "b200phi_p2p_free",
    "b200phi_p2p_status",
    "b200phi_sharded_reduce_p2p",
    "b200phi_last_error",
    "b200phi_version",
]

_lib = None


def lib():
    """Load libb200phi.so once. Raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C paddle2_b200/csrc`. There is no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    L.b200phi_last_error.restype = C.c_char_p
    L.b200phi_version.restype = C.c_char_p
    L.b200phi_reduce_workspace.restype = C.c_size_t
    L.b200phi_reduce_workspace.argtypes = [C.POINTER(ReduceDesc)]
    L.b200phi_reduce.argtypes = [C.POINTER(ReduceDesc), C.c_void_p]
    L.b200phi_reduce_plan_kind.argtypes = [C.POINTER(ReduceDesc)]
    L.b200phi_reduce_plan_kind.restype = C.c_int
    L.b200phi_reduce_noalloc.argtypes = [C.POINTER(ReduceDesc), C.c_void_p]
    L.b200phi_reduce_noalloc.restype = C.c_int
    L.b200phi_argreduce.argtypes = [C.POINTER(ArgReduceDesc), C.c_void_p]
    L.b200phi_argreduce_workspace.restype = C.c_size_t
    L.b200phi_argreduce_workspace.argtypes = [C.POINTER(ArgReduceDesc)]
    L.b200phi_rowstats.argtypes = [C.POINTER(RowStatsDesc), C.c_void_p]
    L.b200phi_elementwise.argtypes = [C.POINTER(ElementwiseDesc), C.c_void_p]
    L.b200phi_map.argtypes = [C.POINTER(MapDesc), C.c_void_p]
    L.b200phi_map.restype = C.c_int
    L.b200phi_logsumexp.argtypes = [C.POINTER(LogsumexpDesc), C.c_void_p]
    L.b200phi_logsumexp.restype = C.c_int
    L.b200phi_mixed_add.argtypes = [C.POINTER(MixedAddDesc), C.c_void_p]
    L.b200phi_mixed_add.restype = C.c_int
    L.b200phi_comm_unique_id.argtypes = [C.c_char_p]
    L.b200phi_comm_init_rank.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    L.b200phi_comm_destroy.argtypes = [C.c_void_p]
    L.b200phi_sharded_reduce.argtypes = [C.POINTER(ReduceDesc), C.c_int64, C.c_void_p, C.c_void_p]
    L.b200phi_sharded_reduce_workspace.restype = C.c_size_t
    L.b200phi_sharded_reduce_workspace.argtypes = [C.POINTER(ReduceDesc)]
    L.b200phi_p2p_status.argtypes = [C.c_void_p]
    L.b200phi_p2p_status.restype = C.c_int
    L.b200phi_sharded_argreduce_workspace.restype = C.c_size_t
    L.b200phi_sharded_argreduce_workspace.argtypes = [C.POINTER(ArgReduceDesc), C.c_int]
    L.b200phi_sharded_argreduce.restype = C.c_int
    L.b200phi_sharded_argreduce.argtypes = [C.POINTER(ArgReduceDesc), C.c_int64, C.c_int, C.c_void_p,
                                            C.c_void_p, C.c_size_t, C.c_void_p]
    L.b200phi_p2p_buffer_bytes.restype = C.c_size_t
    L.b200phi_p2p_buffer_bytes.argtypes = [C.c_int, C.c_int64, C.c_int64]
    L.b200phi_p2p_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]
    L.b200phi_p2p_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.b200phi_p2p_close.argtypes = [C.c_void_p]
    L.b200phi_p2p_free.argtypes = [C.c_void_p]
    L.b200phi_sharded_reduce_p2p.argtypes = [C.POINTER(ReduceDesc), C.c_int64, C.POINTER(P2PGroup),
                                             C.c_uint32, C.c_void_p]
    for name in ("b200phi_p2p_alloc", "b200phi_p2p_open", "b200phi_p2p_close", "b200phi_p2p_free",
                 "b200phi_sharded_reduce_p2p"):
        getattr(L, name).restype = C.c_int
    for name in ("b200phi_reduce", "b200phi_argreduce", "b200phi_rowstats", "b200phi_elementwise",
                 "b200phi_comm_unique_id", "b200phi_comm_init_rank", "b200phi_comm_destroy",
                 "b200phi_sharded_reduce"):
        getattr(L, name).restype = C.c_int
    _lib = L
    return L


def check(status):
    if status != 0:
        raise B200PhiError(status, lib().b200phi_last_error().decode())
