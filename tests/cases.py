"""The parity replay list: shapes/attrs the reference's own tests pin (SURVEY.md
Appendix A), shared by the oracle tests (CPU), the golden-vector generator and the
GPU parity tests.  Source lines are in /root/reference/test/legacy_test/."""
import numpy as np

# (op, dtype, shape, axis, keepdim, source)
REDUCE_CASES = [
    ("sum", "float64", (5, 6, 10), [0], False, "test_reduce_op.py:47-50"),
    ("sum", "float32", (5, 6, 10), [0], False, "test_reduce_op.py:189-261 (fp class factory)"),
    ("sum", "float64", (), [], False, "test_reduce_op.py:87-90"),
    ("sum", "float64", (1, 2, 5, 6, 10), [0], False, "test_reduce_op.py:107-110"),
    ("sum", "float64", (1, 1, 2, 5, 6, 10), [0], False, "test_reduce_op.py:115-118"),
    ("sum", "float64", (1, 3, 1, 2, 1, 4, 3, 10), (0, 3), False, "test_reduce_op.py:123-126"),
    ("sum", "float64", (10, 10), (0, 1), False, "test_reduce_op.py:139-142"),
    ("sum", "float64", (5, 6, 10), (0, 1, 2), False, "test_reduce_op.py:165-168"),
    ("max", "float64", (5, 6, 10), [-1], False, "test_reduce_op.py:276-277"),
    ("max", "float64", (), [], False, "test_reduce_op.py:311"),
    ("max", "float64", (5,), [0], False, "test_reduce_op.py:325"),
    ("max", "float64", (5, 20), [0, 1], False, "test_reduce_op.py:341"),
    ("max", "float32", (5, 6, 10), [-1], True, "test_reduce_op.py:356-361"),
    ("max", "float64", (5, 6, 10), [-2, -1], False, "test_reduce_op.py:1342-1343"),
    ("min", "float64", (5, 6, 10), [1, 2], False, "test_reduce_op.py:1372-1373"),
    ("min", "float64", (5, 6, 10), [2], False, "test_reduce_op.py:431-432"),
    ("min", "float64", (2, 4, 3, 5, 6, 10), [2, 4], False, "test_reduce_op.py:464-466"),
    ("min", "float64", (2, 4, 3, 5, 6, 3, 2, 4), [2, 3, 4], False, "test_reduce_op.py:482-484"),
    ("min", "float32", (5, 6, 10), [2], True, "test_reduce_op.py:509-514"),
    ("prod", "float32", (5, 6, 10), [0], False, "test_reduce_op.py:557"),
    ("prod", "float64", (5, 6, 10), [0], False, "test_reduce_op.py:609"),
    ("prod", "float64", (), [], False, "test_reduce_op.py:647"),
    ("prod", "float64", (100,), None, False, "test_reduce_op.py:660"),
    ("prod", "float64", (5, 6, 10), None, False, "test_reduce_op.py:680"),
    ("prod", "float32", (5, 6, 2, 3, 4, 2), [2, 3, 4], False, "test_reduce_op.py:700-702"),
    ("prod", "float32", (2, 5, 3, 2, 2, 3, 4, 2), [2, 3, 4], False, "test_reduce_op.py:772-776"),
    ("all", "bool", (5, 6, 10), None, False, "test_reduce_op.py:843-869"),
    ("all", "float32", (5, 6, 10), None, False, "test_reduce_op.py:843-869"),
    ("all", "int32", (5, 6, 10), None, False, "test_reduce_op.py:843-869"),
    ("all", "bool", (), [], False, "test_reduce_op.py:879-881"),
    ("all", "bool", (2, 5, 3, 2, 2, 3, 4, 2), (2, 3, 4), False, "test_reduce_op.py:892"),
    ("all", "bool", (2, 5, 3, 2, 2, 3, 4, 2), (1, 3, 4), False, "test_reduce_op.py:920"),
    ("all", "bool", (2, 5, 3, 2, 2, 3, 4, 2), (5,), True, "test_reduce_op.py:954"),
    ("all", "bool", (5, 6, 10), (1,), False, "test_reduce_op.py:907-908"),
    ("all", "bool", (5, 6, 10), [1], True, "test_reduce_op.py:935-936"),
    ("any", "bool", (5, 6, 10), None, False, "test_reduce_op.py:986-1016"),
    ("any", "float32", (5, 6, 10), None, False, "test_reduce_op.py:986-1016"),
    ("any", "int64", (5, 6, 10), None, False, "test_reduce_op.py:986-1016"),
    ("any", "bool", (), [], False, "test_reduce_op.py:1028-1030"),
    ("any", "bool", (2, 5, 3, 2, 2, 3, 4, 2), (3, 5, 4), False, "test_reduce_op.py:1043"),
    ("any", "bool", (2, 5, 3, 2, 2, 3, 4, 2), (3, 6), False, "test_reduce_op.py:1075"),
    ("any", "bool", (2, 5, 3, 2, 2, 3, 4, 2), (1,), True, "test_reduce_op.py:1115"),
    ("any", "bool", (5, 6, 10), [1], False, "test_reduce_op.py:1060-1061"),
    ("any", "bool", (5, 6, 10), (1,), True, "test_reduce_op.py:1092-1093"),
    ("sum", "float64", (120,), None, False, "test_reduce_op.py:1143"),
    ("sum", "float64", (20, 10), [0], False, "test_reduce_op.py:1174"),
    ("sum", "float64", (20, 10), [1], False, "test_reduce_op.py:1187"),
    ("sum", "float64", (5, 6, 7), [1], False, "test_reduce_op.py:1200"),
    ("sum", "float64", (5, 6, 7), [2], False, "test_reduce_op.py:1213"),
    ("sum", "float64", (5, 6, 7), [-2], False, "test_reduce_op.py:1226"),
    ("sum", "float64", (5, 6, 7), [1, 2], False, "test_reduce_op.py:1243"),
    ("sum", "float64", (2, 5, 3, 2, 2, 3, 4, 2), (4, 2, 3), False, "test_reduce_op.py:1262-1264"),
    ("sum", "float64", (2, 5, 3, 2, 2, 3, 4, 2), (3, 4, 5), True, "test_reduce_op.py:1314-1316"),
    ("sum", "float64", (5, 6, 10), [1], True, "test_reduce_op.py:1283-1284"),
    ("sum", "float64", (5, 6, 10), [-2, -1], True, "test_reduce_op.py:1388-1389"),
    ("sum", "float64", (100, 1, 1), [1, 2], True, "test_reduce_op.py:1432-1433"),
    ("sum", "float64", (100, 1), [1], False, "test_reduce_op.py:1477-1478"),
    ("sum", "float64", (100, 1, 1), None, False, "test_reduce_op.py:1508"),
    ("sum", "float32", (100, 1, 1), None, False, "test_reduce_op.py:1530"),
    ("sum", "float64", (100,), [0], False, "test_reduce_op.py:1550-1551"),
    ("mean", "float32", (5, 6, 10), [1], False, "test_mean_op.py"),
    ("mean", "float64", (5, 6, 10), None, False, "test_mean_op.py"),
    ("mean", "float32", (3, 4, 5), [0, 2], True, "test_mean_op.py"),
    ("amax", "float32", (5, 6, 10), [-1], False, "test_max_op.py / amax"),
    ("amin", "float64", (5, 6, 10), [0], False, "test_min_op.py / amin"),
    ("max", "int32", (5, 6, 10), [1], False, "cpu/reduce_max_kernel.cc:38-39 dtype list"),
    ("min", "int64", (5, 6, 10), [0, 2], False, "cpu/reduce_min_kernel.cc dtype list"),
    ("prod", "int32", (3, 4, 5), [1], False, "cpu/prod_kernel.cc:39-40 dtype list"),
    ("prod", "int64", (3, 4, 5), None, False, "cpu/prod_kernel.cc:39-40 dtype list"),
]

# (in dtype, out dtype, shape, axis, keepdim, source): sum with an explicit / promoted dtype
SUM_DTYPE_CASES = [
    ("float32", "float64", (6, 2, 10), None, False, "test_reduce_op.py:1577-1598"),
    ("float32", "float64", (6, 2, 10), [1], False, "test_reduce_op.py:1610"),
    ("float32", "float64", (6, 2, 10), [1], True, "test_reduce_op.py:1632"),
    ("int32", "int64", (2, 3, 4), None, False, "test_reduce_op.py:1731 (int32 -> int64 promotion)"),
    ("bool", "int64", (4, 4), [0], False, "infermeta/unary.cc:4698-4707"),
    ("int64", "int64", (12, 10), [1], False, "reduce_sum_kernel.cc:40-57"),
    ("int64", "float32", (7, 9), [0], False, "gpu/reduce.h:48-69 cast-then-reduce"),
    ("float64", "float32", (7, 9), [1], False, "gpu/reduce.h:48-69 cast-then-reduce"),
    ("uint8", "uint8", (5, 300), [1], False, "reduce_sum_kernel.cc:40-57 (wraps mod 256)"),
    ("int16", "int16", (5, 300), [1], False, "reduce_sum_kernel.cc:40-57"),
    ("int8", "int64", (5, 300), [0], False, "cast-then-reduce"),
]

# (op, dtype, shape, axis, keepdim, index dtype, source)
ARG_CASES = [
    ("argmin", "float32", (3, 4, 5), 0, False, "int64", "test_arg_min_max_op.py:30-50"),
    ("argmax", "float32", (3, 4, 5), 0, False, "int64", "test_arg_min_max_op.py:30-50"),
    ("argmax", "float64", (3, 4), 1, False, "int64", "test_arg_min_max_op.py:53-76"),
    ("argmin", "int64", (3, 4), 0, False, "int64", "test_arg_min_max_op.py:53-76"),
    ("argmax", "float32", (4, 5, 6), 0, False, "int64", "test_arg_min_max_v2_op.py:31"),
    ("argmax", "float32", (4, 5, 6), 1, False, "int64", "test_arg_min_max_v2_op.py"),
    ("argmax", "float32", (4, 5, 6), 2, False, "int64", "test_arg_min_max_v2_op.py"),
    ("argmin", "float32", (4, 5, 6), -1, False, "int64", "test_arg_min_max_v2_op.py"),
    ("argmin", "float32", (4, 5, 6), -2, False, "int64", "test_arg_min_max_v2_op.py"),
    ("argmax", "float32", (4, 5, 6), 0, True, "int64", "test_arg_min_max_v2_op.py keepdims"),
    ("argmax", "float32", (4, 5, 6), None, False, "int64", "test_arg_min_max_v2_op.py flatten"),
    ("argmin", "float32", (4, 5, 6), None, True, "int64", "test_arg_min_max_v2_op.py flatten+keepdims"),
    ("argmax", "float32", (4, 5, 6), 1, False, "int32", "test_arg_min_max_v2_op.py dtype=int32"),
    ("argmax", "float32", (5, 5, 5, 5), 0, False, "int64", "test_arg_min_max_v2_op.py:261-293"),
    ("argmax", "float32", (4, 4, 4, 4, 4), 0, False, "int64", "test_arg_min_max_v2_op.py:261-293"),
    ("argmin", "float32", (3, 3, 3, 3, 3, 3), 0, False, "int64", "test_arg_min_max_v2_op.py:261-293"),
    ("argmax", "int32", (7, 33), 1, False, "int64", "cpu/arg_min_max_kernel.cc:200-224 dtype list"),
    ("argmax", "float32", (), None, False, "int64", "0-D input => 0"),
]

# (x shape, y shape, axis, source): broadcast add
ADD_CASES = [
    ((13, 17), (13, 17), -1, "test_elementwise_add_op.py:109-110"),
    ((), (), -1, "test_elementwise_add_op.py:128"),
    ((), (13, 17), -1, "test_elementwise_add_op.py:135"),
    ((13, 17), (), -1, "test_elementwise_add_op.py:143"),
    ((2, 3, 4), (1,), -1, "test_elementwise_add_op.py:269-270"),
    ((2, 3, 4), (1, 1), -1, "test_elementwise_add_op.py:289-290"),
    ((100,), (100,), -1, "test_elementwise_add_op.py:306-307"),
    ((100, 2, 3), (100,), 0, "test_elementwise_add_op.py:320-326"),
    ((2, 100, 3), (100,), 1, "test_elementwise_add_op.py:364-370"),
    ((2, 3, 100), (100,), -1, "test_elementwise_add_op.py:391-393"),
    ((2, 10, 12, 1), (10, 12), 1, "test_elementwise_add_op.py:412-418"),
    ((2, 10, 12, 3), (10, 12), 1, "test_elementwise_add_op.py:425-431"),
    ((100, 2, 1, 2), (100, 1), 0, "test_elementwise_add_op.py:436-442"),
    ((10, 3, 12), (10, 1, 12), -1, "test_elementwise_add_op.py:460-461"),
    ((2, 12, 3, 5), (2, 12, 1, 5), -1, "test_elementwise_add_op.py:476-477"),
    ((1, 1, 20, 5), (20, 5, 1, 1), -1, "test_elementwise_add_op.py:483-484"),
    ((2, 10, 12), (10, 12), 1, "test_elementwise_add_op.py:499-504"),
    ((10, 100, 1), (100, 1), -1, "test_elementwise_add_op.py:533-534"),
    ((100, 1), (1,), -1, "test_elementwise_add_op.py:543-544"),
    ((100, 2, 3), (100, 1, 1), -1, "test_elementwise_add_op.py:550-555"),
    ((2, 3, 100), (1, 1, 100), -1, "test_elementwise_add_op.py:570-575"),
    ((10, 3, 1, 4), (10, 1, 12, 1), -1, "test_elementwise_add_op.py:590-595"),
    ((10, 12), (2, 2, 10, 12), 2, "test_elementwise_add_op.py:600-605"),
    ((10, 1, 12), (10, 2, 12), 0, "test_elementwise_add_op.py:610-615"),
]

# (x shape, y shape, axis, source): fused_elemwise_activation shapes
FUSED_SHAPES = [
    ((13, 17), (13, 17), -1, "test_fused_elemwise_activation_op.py:85-86"),
    ((2, 3, 4), (1,), -1, "test_fused_elemwise_activation_op.py:162-163"),
    ((2, 3, 4), (1, 1), -1, "test_fused_elemwise_activation_op.py:169-170"),
    ((32,), (32,), -1, "test_fused_elemwise_activation_op.py:176-177"),
    ((2, 3, 4), (2,), 0, "test_fused_elemwise_activation_op.py:183-185"),
    ((2, 3, 4), (3,), 1, "test_fused_elemwise_activation_op.py:196-198"),
    ((2, 3, 4), (4,), -1, "test_fused_elemwise_activation_op.py:209-210"),
    ((2, 3, 4, 5), (3, 4), 1, "test_fused_elemwise_activation_op.py:221-223"),
    ((2, 3, 4, 5), (2, 1), 0, "test_fused_elemwise_activation_op.py:234-236"),
    ((2, 3, 4), (3, 4), 1, "test_fused_elemwise_activation_op.py:247-249"),
    ((2, 1), (1,), 1, "test_fused_elemwise_activation_op.py:260-262"),
    ((3, 20, 20), (3, 1, 1), -1, "test_fused_elemwise_activation_op.py:273-274"),
]
FUSED_FUNCTOR_LISTS = [
    ["scale", "elementwise_add"], ["elementwise_add", "scale"],
    ["relu", "elementwise_add"], ["elementwise_add", "relu"],
    ["elementwise_mul", "scale"], ["gelu", "elementwise_add"],
]


def make_input(rng, dtype, shape, kind="uniform"):
    """Inputs the way the reference tests draw them (np.random.random, seed 123,
    op_test.py:437-438); integers 0..9, bools ~50/50."""
    dt = np.dtype(dtype)
    if dt == np.bool_:
        return rng.integers(0, 2, size=shape).astype(np.bool_)
    if np.issubdtype(dt, np.integer):
        hi = 4 if kind == "prod" else 10
        return rng.integers(0, hi, size=shape).astype(dt)
    if kind == "logical":
        return rng.integers(0, 2, size=shape).astype(dt)
    if kind == "signed":
        return (rng.random(shape) * 2 - 1).astype(dt)
    return rng.random(shape).astype(dt)


# ---- functor plug-ins (b200phi_map) ------------------------------------------------------
# (x_shape, y_shape, source): the broadcast matrix of test_logical_op.py:45-60, which
# test_bitwise_op.py:57-116 and test_compare_op.py:317-465 repeat in part
MAP_SHAPES = [
    ([2, 3, 4, 5], [4, 5], "test_logical_op.py:45"),
    ([2, 3, 4, 5], [4, 1], "test_logical_op.py:46"),
    ([2, 3, 4, 5], [1, 4, 1], "test_logical_op.py:47"),
    ([2, 3, 4, 5], [3, 4, 1], "test_logical_op.py:48"),
    ([2, 3, 1, 5], [3, 1, 1], "test_logical_op.py:49"),
    ([4, 1], [2, 3, 4, 5], "test_logical_op.py:50"),
    ([1, 4, 1], [2, 3, 4, 5], "test_logical_op.py:51"),
    ([3, 4, 1], [2, 3, 4, 5], "test_logical_op.py:52"),
    ([3, 1, 1], [2, 3, 1, 5], "test_logical_op.py:53"),
    ([4, 5], [2, 3, 4, 5], "test_logical_op.py:54"),
    ([1, 4, 5], [2, 3, 1, 5], "test_logical_op.py:55"),
    ([10, 7], [10, 7], "test_logical_op.py:56"),
    ([1, 1, 4, 5], [2, 3, 1, 5], "test_logical_op.py:57"),
    ([], [], "test_logical_op.py:58"),
    ([2, 3, 4, 5], [], "test_logical_op.py:59"),
    ([], [2, 3, 4, 5], "test_logical_op.py:60"),
    ([1, 2, 1, 3], [1, 2, 3], "test_compare_op.py:317-319"),
    ([5], [3, 1], "test_compare_op.py:356-357"),
    ([3, 1], [1], "test_compare_op.py:446-447"),
]
MAP_ERROR_SHAPES = [
    ([2, 3, 4, 5], [3, 4], "test_logical_op.py:64"),
    ([2, 3, 4, 5], [4, 3], "test_logical_op.py:65"),
]
COMPARE_OPS = ["equal", "not_equal", "less_than", "less_equal", "greater_than", "greater_equal"]
LOGICAL_OPS = ["logical_and", "logical_or", "logical_xor"]
BITWISE_OPS = ["bitwise_and", "bitwise_or", "bitwise_xor"]
SHIFT_OPS = ["left_shift_arithmetic", "left_shift_logic", "right_shift_arithmetic", "right_shift_logic"]
SHIFT_DTYPES = ["uint8", "int8", "int16", "int32", "int64"]
MAP_INT_DTYPES = ["bool", "uint8", "int8", "int16", "int32", "int64"]
MAP_ALL_DTYPES = MAP_INT_DTYPES + ["float32", "float64"]

# (name, p0, p1, domain): attributes are the defaults of python/paddle/nn/functional/activation.py
# or the values test_activation_op.py sets; domain "pos" keeps log/sqrt arguments positive
ACTIVATIONS = [
    ("exp", 0, 0, "any"), ("log", 0, 0, "pos"), ("sqrt", 0, 0, "pos"), ("rsqrt", 0, 0, "pos"),
    ("square", 0, 0, "any"), ("reciprocal", 0, 0, "pos"), ("sin", 0, 0, "any"),
    ("cos", 0, 0, "any"), ("tanh", 0, 0, "any"), ("sigmoid", 0, 0, "any"), ("silu", 0, 0, "any"),
    ("leaky_relu", 0.02, 0, "any"), ("elu", 1.0, 0, "any"), ("softplus", 2.0, 15.0, "any"),
    ("floor", 0, 0, "any"), ("ceil", 0, 0, "any"), ("expm1", 0, 0, "any"), ("log1p", 0, 0, "pos"),
    ("log2", 0, 0, "pos"), ("log10", 0, 0, "pos"), ("relu6", 6.0, 0, "wide"),
    ("hard_sigmoid", 0.2, 0.5, "wide"), ("swish", 1.0, 0, "any"), ("mish", 20.0, 0, "any"),
    ("softsign", 0, 0, "any"), ("log_sigmoid", 0, 0, "any"), ("abs", 0, 0, "any"),
    ("pow_abs", 2.5, 0, "any"), ("relu", 0, 0, "any"),
]
ACT_SHAPES = [[11, 17], [], [10, 12], [3, 1000], [4097]]  # test_activation_op.py:96,548,5283 + vector paths


def map_operand(rng, dtype, shape, domain="any"):
    dt = np.dtype(dtype)
    if dt == np.bool_:
        return rng.integers(0, 2, size=shape).astype(np.bool_)
    if dt == np.uint8:
        return rng.integers(0, 100, size=shape).astype(dt)
    if np.issubdtype(dt, np.integer):
        return rng.integers(-100, 100, size=shape).astype(dt)  # test_bitwise_op.py:61-62
    if domain == "pos":
        return (rng.random(shape) * 3 + 0.1).astype(dt)
    if domain == "wide":
        return (rng.random(shape) * 16 - 8).astype(dt)
    # few distinct values so that compare ops see ties
    return (np.round(rng.random(shape) * 8 - 4, 1)).astype(dt) if domain == "ties" \
        else (rng.random(shape) * 4 - 2).astype(dt)


# ---- reduce-side plug-ins ----------------------------------------------------------------
# (shape, axis, porder, keepdim, asvector, source): test_norm_all.py's p_norm cases + every porder branch
PNORM_CASES = [
    ([2, 3, 4, 5], 1, 2.0, False, False, "test_norm_all.py:217-222"),
    ([3, 20, 3], 2, 2.0, True, False, "test_norm_all.py:266-271"),
    ([3, 20, 3], 2, float("inf"), True, False, "test_norm_all.py:282-287"),
    ([3, 20, 3], 2, float("-inf"), True, False, "test_norm_all.py:300-305"),
    ([3, 20, 3], 2, 0.0, True, False, "test_norm_all.py:318-323"),
    ([3, 20, 3], -1, 2.0, False, False, "test_norm_all.py:334-339"),
    ([3, 20, 3], 1, 1.0, False, False, "porder 1"),
    ([3, 20, 3], 0, 3.0, False, False, "general porder"),
    ([3, 20, 3], 1, 0.5, True, False, "fractional porder"),
    ([2, 3, 4, 5], 0, 2.0, False, True, "asvector"),
    ([64, 1030], -1, 2.0, False, False, "long rows"),
    ([1030, 64], 0, 1.0, False, False, "columns"),
]
# (shape, axis, keepdim, source): test_logsumexp.py
LOGSUMEXP_CASES = [
    ([2, 3, 4, 5], [-1], False, "test_logsumexp.py:63-65"),
    ([], [], False, "test_logsumexp.py:116-117"),
    ([4, 5, 6], [-1], False, "test_logsumexp.py:122"),
    ([2, 3, 4, 5], [0, -1], False, "test_logsumexp.py:127"),
    ([2, 3, 4, 5], [0, 1, 2, 3], False, "test_logsumexp.py:132"),
    ([2, 3, 4, 5], [-1], True, "keepdim"),
    ([2, 3, 4, 5], [2, 3], False, "two trailing axes"),
    ([2, 3, 4, 5], [1], False, "middle axis: composed route"),
    ([2, 3, 4, 5], [0], True, "leading axis: composed route"),
    ([37, 1000], [-1], False, "vocabulary-sized rows"),
    ([5, 4099], [-1], False, "ragged rows (scalar route)"),
]


# ---- the arithmetic binary functors registered by kps/elementwise_kernel.cu:183-293 next to add ..
# minimum: (name, dtypes as the reference registers them)
ARITH_OPS = [
    ("remainder", ["float32", "float64", "int32", "int64"]),
    ("floor_divide", ["uint8", "int8", "int16", "int32", "int64", "float32", "float64"]),
    ("elementwise_pow", ["float32", "float64", "int32", "int64"]),
    ("copysign", ["bool", "uint8", "int8", "int16", "int32", "int64", "float32", "float64"]),
    ("fmax", ["float32", "float64", "int32", "int64"]),
    ("fmin", ["float32", "float64", "int32", "int64"]),
    ("heaviside", ["float32", "float64", "int32", "int64"]),
]
ARITH_SHAPES = [([7, 33], [7, 33]), ([5, 1, 9], [4, 1]), ([3, 4], [1]), ([2, 3, 4, 5], [3, 1, 5]), ([40], [6, 40])]


def arith_operands(rng, name, dtype, xs, ys):
    """Operands inside each functor's domain: no integer division by zero (the reference asserts
    there), small non-negative integer exponents, positive float bases for pow; zeros, sign
    changes and exact ties elsewhere."""
    dt = np.dtype(dtype)
    if dt == np.bool_:
        return rng.integers(0, 2, size=xs).astype(dt), rng.integers(0, 2, size=ys).astype(dt)
    if np.issubdtype(dt, np.integer):
        lo = 0 if dt == np.uint8 else -9
        x = rng.integers(lo, 10, size=xs).astype(dt)
        y = rng.integers(lo, 10, size=ys).astype(dt)
        if name in ("remainder", "floor_divide"):
            y[y == 0] = 3
        if name == "elementwise_pow":
            y = (np.abs(y) % 5).astype(dt)
        return x, y
    x = np.round(rng.standard_normal(xs) * 5, 1).astype(dt)
    y = np.round(rng.standard_normal(ys) * 3, 1).astype(dt)
    x.reshape(-1)[::7] = 0
    if name == "elementwise_pow":
        x = (np.abs(x) + 0.1).astype(dt)
    if name in ("remainder", "floor_divide"):
        y[y == 0] = dt.type(1.5)
    return x, y
