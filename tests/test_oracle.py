"""Pins the oracle (oracle/phi_oracle.c, the plain-C port) before anything trusts it:
  * against the committed golden vectors that the REFERENCE build produced
    (tests/golden/phi_ref_golden.npz, generator tests/golden/make_golden.py),
  * against the reference build itself when oracle/_ref/libphi_ref.so is present
    (it is built here from /root/reference and travels to the GPU box),
  * against the literal answers in the reference's docstrings.
Bit-exact for every dtype, floating point included: the port restates Eigen's
accumulation order."""
import os

import numpy as np
import pytest

import cases

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_ref_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("i", range(len(cases.REDUCE_CASES)))
def test_reduce_port_matches_golden(oracle, golden, i):
    op, dt, shape, axis, keepdim, src = cases.REDUCE_CASES[i]
    x = golden[f"reduce_{i}_x"]
    want = golden[f"reduce_{i}_y"]
    got = oracle.reduce(op, x, axis, keepdim)
    assert got.shape == want.shape and got.dtype == want.dtype, src
    assert np.array_equal(got, want), src


@pytest.mark.parametrize("i", range(len(cases.SUM_DTYPE_CASES)))
def test_sum_dtype_port(oracle, golden, i):
    idt, odt, shape, axis, keepdim, src = cases.SUM_DTYPE_CASES[i]
    x = golden[f"sumdt_{i}_x"]
    got = oracle.reduce("sum", x, axis, keepdim, out_dtype=odt)
    if f"sumdt_{i}_y" in golden.files:
        assert np.array_equal(got, golden[f"sumdt_{i}_y"]), src
    else:  # narrow integer outputs: the reference casts, then accumulates in that type
        want = x.astype(odt).sum(axis=tuple(axis) if axis is not None else None, dtype=odt,
                                 keepdims=keepdim)
        assert np.array_equal(got, want), src


@pytest.mark.parametrize("i", range(len(cases.ARG_CASES)))
def test_arg_port_matches_golden(oracle, golden, i):
    op, dt, shape, axis, keepdim, idt, src = cases.ARG_CASES[i]
    x = golden[f"arg_{i}_x"]
    want = golden[f"arg_{i}_y"]
    got = oracle.argreduce(op, x, axis, keepdim, np.dtype(idt))
    assert got.dtype == want.dtype and got.shape == want.shape, src
    assert np.array_equal(got, want), src
    # and numpy agrees (the reference tests' own oracle, test_arg_min_max_op.py:40-50)
    npf = np.argmax if op == "argmax" else np.argmin
    ref = npf(x) if axis is None else npf(x, axis=axis)
    assert np.array_equal(got.reshape(np.shape(ref)), ref), src


@pytest.mark.parametrize("i", range(len(cases.ADD_CASES)))
def test_add_port_matches_golden(oracle, golden, i):
    xs, ys, axis, src = cases.ADD_CASES[i]
    got = oracle.elementwise("add", [golden[f"add_{i}_x"], golden[f"add_{i}_y"]], axis=axis)
    assert np.array_equal(got, golden[f"add_{i}_z"]), src


def test_fused_port_matches_golden(oracle, golden):
    from golden.make_golden import parse_functor_list
    k = 0
    for (xs, ys, axis, src) in cases.FUSED_SHAPES:
        for fl in cases.FUSED_FUNCTOR_LISTS:
            binary, act, compound = parse_functor_list(fl)
            out, inter = oracle.elementwise(binary, [golden[f"fused_{k}_x"], golden[f"fused_{k}_y"]],
                                            act=act, compound=compound, scale=0.1, axis=axis,
                                            want_intermediate=True)
            np.testing.assert_allclose(out, golden[f"fused_{k}_out"], rtol=1e-6, atol=1e-7,
                                       err_msg=f"{src} {fl}")
            np.testing.assert_allclose(inter, golden[f"fused_{k}_inter"], rtol=1e-6, atol=1e-7)
            k += 1


def test_ternary_and_var_port(oracle, golden):
    out = oracle.elementwise("add", [golden["ternary_a"], golden["ternary_b"], golden["ternary_c"]])
    assert np.array_equal(out, golden["ternary_out"])
    mean, var = oracle.rowstats_f32(golden["var_x"], unbiased=True)
    assert np.array_equal(var, golden["var_y"])
    assert np.array_equal(mean, golden["mean_y"])
    # the reference test's own oracle: numpy variance with ddof=1 (test_variance_layer.py)
    np.testing.assert_allclose(var, np.var(golden["var_x"].astype(np.float64), axis=-1, ddof=1),
                               rtol=1e-4)


def test_docstring_examples(oracle):
    """Literal answers in python/paddle/tensor/math.py (sum :1475-1567, max :2870-2987)
    and search.py (argmax)."""
    x = np.array([[0.2, 0.3, 0.5, 0.9], [0.1, 0.2, 0.6, 0.7]], dtype=np.float32)
    np.testing.assert_allclose(oracle.reduce("sum", x, None), 3.5, rtol=1e-6)
    np.testing.assert_allclose(oracle.reduce("sum", x, [0]), [0.3, 0.5, 1.1, 1.6], rtol=1e-6)
    np.testing.assert_allclose(oracle.reduce("sum", x, [-1]), [1.9, 1.6], rtol=1e-6)
    np.testing.assert_allclose(oracle.reduce("sum", x, [-1], keepdim=True), [[1.9], [1.6]], rtol=1e-6)
    y = np.array([[[1, 2], [3, 4]], [[5, 6], [7, 8]]], dtype=np.float32)
    np.testing.assert_allclose(oracle.reduce("sum", y, [1, 2]), [10, 26])
    np.testing.assert_allclose(oracle.reduce("sum", y, [0, 1]), [16, 20])
    b = np.array([[True, True, True, True], [False, False, False, False]])
    assert oracle.reduce("sum", b, None, out_dtype=np.int64) == 4
    assert np.array_equal(oracle.reduce("sum", b, [0], out_dtype=np.int64), [1, 1, 1, 1])
    assert np.array_equal(oracle.reduce("sum", b, [1], out_dtype=np.int64), [4, 0])
    np.testing.assert_allclose(oracle.reduce("max", x, None), 0.9)
    np.testing.assert_allclose(oracle.reduce("max", x, [0]), [0.2, 0.3, 0.6, 0.9])
    np.testing.assert_allclose(oracle.reduce("max", x, [-1]), [0.9, 0.7])
    a = np.array([[5, 8, 9, 5], [0, 0, 1, 7], [6, 9, 2, 4]], dtype=np.float32)
    assert oracle.argreduce("argmax", a, None) == 2
    assert np.array_equal(oracle.argreduce("argmax", a, 0), [2, 2, 0, 1])
    assert np.array_equal(oracle.argreduce("argmax", a, -1), [2, 3, 1])
    assert np.array_equal(oracle.argreduce("argmin", a, -1), [0, 0, 2])


def test_tie_break_and_identity_rows(oracle):
    """SURVEY.md probe P7: lowest index on ties; an all-lowest() row answers 0."""
    i64 = np.iinfo(np.int64)
    x = np.array([[1, 3, 3, 2], [i64.min] * 4, [5, 1, 5, 5]], dtype=np.int64)
    assert np.array_equal(oracle.argreduce("argmax", x, 1), [1, 0, 0])
    y = np.array([[i64.max] * 4, [2, 1, 1, 3]], dtype=np.int64)
    assert np.array_equal(oracle.argreduce("argmin", y, 1), [0, 1])


def test_errors(oracle):
    with pytest.raises(ValueError):
        oracle.reduce("sum", np.zeros((0, 3), np.float32), [0])  # "must not empty"
    with pytest.raises(ValueError):
        oracle.elementwise("add", [np.zeros((2, 3), np.float32), np.zeros((2, 4), np.float32)])


# ---- against the reference build itself, on fresh random inputs -----------------------------
def _need_ref(oracle):
    if not oracle.have_ref():
        pytest.skip("oracle/_ref/libphi_ref.so not built (needs /root/reference)")


@pytest.mark.parametrize("shape,axis", [
    ((5, 6, 10), [0]), ((5, 6, 10), [-1]), ((32, 4096), [-1]), ((3, 100000), [1]),
    ((100000,), None), ((2, 4, 3, 5, 6, 10), [2, 4]), ((2, 5, 3, 2, 2, 3, 4, 2), (1, 3, 4)),
    ((5, 6, 10), [0, 2]), ((20000, 16), [0]), ((7, 33, 9), [1]), ((100, 1, 1), [1, 2]),
])
def test_port_equals_reference_build_reduce(oracle, shape, axis):
    _need_ref(oracle)
    rng = np.random.default_rng(123)
    for dt in (np.float32, np.float64, np.int32, np.int64):
        x = (rng.random(shape) * (10 if np.issubdtype(dt, np.integer) else 1)).astype(dt)
        for op in ("sum", "mean", "max", "min", "prod"):
            if op == "mean" and np.issubdtype(dt, np.integer):
                continue
            a = oracle.reduce(op, x, axis, impl="port")
            b = oracle.reduce(op, x, axis, impl="ref")
            assert np.array_equal(a, b), (shape, axis, dt, op)
    xb = rng.integers(0, 2, size=shape).astype(np.bool_)
    for op in ("any", "all"):
        assert np.array_equal(oracle.reduce(op, xb, axis), oracle.reduce(op, xb, axis, impl="ref"))


def test_port_equals_reference_build_arg_and_elementwise(oracle):
    _need_ref(oracle)
    rng = np.random.default_rng(7)
    for shape, axis in [((3, 4, 5), 0), ((3, 4), 1), ((3, 4, 5), None), ((64, 257), 1)]:
        for dt in (np.float32, np.float64, np.int32, np.int64):
            x = (rng.random(shape) * 4).astype(dt)  # few distinct ints: ties
            for op in ("argmax", "argmin"):
                assert np.array_equal(oracle.argreduce(op, x, axis),
                                      oracle.argreduce(op, x, axis, impl="ref"))
    for xs, ys, ax, _ in cases.ADD_CASES:
        for dt in (np.float32, np.float64, np.int32, np.int64):
            x = (rng.random(xs) * 5 - 2).astype(dt)
            y = (rng.random(ys) * 5 - 2).astype(dt)
            for b in ("add", "subtract", "multiply", "maximum", "minimum"):
                assert np.array_equal(oracle.elementwise(b, [x, y], axis=ax),
                                      oracle.elementwise(b, [x, y], axis=ax, impl="ref"))


# ---- functor plug-ins (b200phi_map) ---------------------------------------------------------
@pytest.mark.parametrize("i", range(len(cases.MAP_SHAPES)))
def test_map_port_matches_golden(oracle, golden, i):
    """compare / logical / bitwise: the port is bit-identical to the reference's functors."""
    _xs, _ys, src = cases.MAP_SHAPES[i]
    for dt in ("float32", "int32", "bool"):
        x, y = golden[f"map_{i}_{dt}_x"], golden[f"map_{i}_{dt}_y"]
        names = cases.COMPARE_OPS + cases.LOGICAL_OPS + ([] if dt == "float32" else cases.BITWISE_OPS)
        for name in names:
            want = golden[f"map_{i}_{dt}_{name}"]
            got = oracle.map(name, [x, y])
            assert got.dtype == want.dtype and got.shape == want.shape, (src, name, dt)
            assert np.array_equal(got, want), (src, name, dt)


def test_map_grad_and_activation_port(oracle, golden):
    x, y, g = golden["mapgrad_x"], golden["mapgrad_y"], golden["mapgrad_g"]
    for name in ("equal_mul", "max_grad_x", "max_grad_y", "min_grad_x", "min_grad_y", "div_grad_y"):
        assert np.array_equal(oracle.map(name, [x, y, g]), golden[f"mapgrad_{name}"]), name
    for name, p0, p1, _dom in cases.ACTIVATIONS:
        got = oracle.map(name, [golden[f"act_{name}_x"]], p0, p1)
        # libm here, Eigen's packet math in the reference: a few ulp apart on transcendentals
        np.testing.assert_allclose(got, golden[f"act_{name}_y"], rtol=2e-6, atol=1e-7, err_msg=name)


def test_map_equal_is_the_tolerant_form(oracle):
    """EqualFunctor on floats (compare_functors.h:34-48): |a-b| < 1e-8, inf by ==, NaN never."""
    a = np.array([1e-9, 0.0, np.inf, np.inf, np.nan, 1.0, 1.0], dtype=np.float32)
    b = np.array([2e-9, 9e-9, np.inf, -np.inf, np.nan, 1.0, np.nextafter(np.float32(1), np.float32(2))],
                 dtype=np.float32)
    assert oracle.map("equal", [a, b]).tolist() == [True, True, True, False, False, True, False]
    assert oracle.map("not_equal", [a, b]).tolist() == [False, False, False, True, True, False, True]
    if oracle.have_ref():
        assert np.array_equal(oracle.map("equal", [a, b], impl="ref"), oracle.map("equal", [a, b]))


def test_map_errors(oracle):
    for xs, ys, _src in cases.MAP_ERROR_SHAPES:
        with pytest.raises(ValueError):
            oracle.map("logical_and", [np.zeros(xs, np.bool_), np.zeros(ys, np.bool_)])
    with pytest.raises(ValueError):
        oracle.map("bitwise_and", [np.zeros(3, np.float32), np.zeros(3, np.float32)])
    with pytest.raises(ValueError):
        oracle.map("exp", [np.zeros(3, np.int32)])


# ---- reduce-side plug-ins ---------------------------------------------------------------------
def test_norm_and_logsumexp_port_match_golden(oracle, golden):
    for i, (shape, axis, porder, keepdim, asvector, src) in enumerate(cases.PNORM_CASES):
        got = oracle.p_norm(golden[f"pnorm_{i}_x"], porder, axis, keepdim, asvector)
        want = golden[f"pnorm_{i}_y"]
        assert got.shape == want.shape, src
        if porder in (0.0, float("inf"), float("-inf")):
            assert np.array_equal(got, want), src
        else:
            np.testing.assert_allclose(got, want, rtol=2e-6, err_msg=src)
    np.testing.assert_allclose(oracle.squared_l2_norm(golden["sql2_x"]), golden["sql2_y"], rtol=1e-6)
    for i, (shape, axis, keepdim, src) in enumerate(cases.LOGSUMEXP_CASES):
        x = golden[f"lse_{i}_x"]
        got = oracle.logsumexp(x, axis if axis else None, keepdim)
        if f"lse_{i}_y" in golden.files:
            assert got.shape == golden[f"lse_{i}_y"].shape, src
            np.testing.assert_allclose(got, golden[f"lse_{i}_y"], rtol=2e-6, atol=1e-6, err_msg=src)
        # the reference test's own oracle (test_logsumexp.py:24-32)
        ax = tuple(axis) if axis else None
        np.testing.assert_allclose(got, np.log(np.exp(x.astype(np.float64)).sum(axis=ax, keepdims=keepdim)),
                                   rtol=1e-5, atol=1e-6, err_msg=src)


def test_mixed_add_port_matches_golden(oracle, golden):
    x = golden["mpadd_x"]
    assert np.array_equal(oracle.mixed_add(x, golden["mpadd_bf16_bits"], True), golden["mpadd_bf16_out"])
    assert np.array_equal(oracle.mixed_add(x, golden["mpadd_fp16_bits"], False), golden["mpadd_fp16_out"])


def test_shift_port_matches_golden(oracle, golden):
    for dt in cases.SHIFT_DTYPES:
        x, y = golden[f"shift_{dt}_x"], golden[f"shift_{dt}_y"]
        for name in cases.SHIFT_OPS:
            assert np.array_equal(oracle.map(name, [x, y]), golden[f"shift_{dt}_{name}"]), (dt, name)


@pytest.mark.parametrize("name,dtypes", cases.ARITH_OPS)
def test_arith_binary_port_matches_golden(oracle, golden, name, dtypes):
    """remainder / floor_divide / elementwise_pow / copysign / fmax / fmin / heaviside: the C port
    against the arrays the reference's own functors produced, bit for bit."""
    for dt in dtypes:
        for j in range(len(cases.ARITH_SHAPES)):
            x, y = golden[f"arith_{name}_{dt}_{j}_x"], golden[f"arith_{name}_{dt}_{j}_y"]
            want = golden[f"arith_{name}_{dt}_{j}_out"]
            got = oracle.map(name, [x, y])
            assert got.dtype == want.dtype and got.shape == want.shape
            if name in ("fmax", "fmin"):   # fmax(+0, -0) may be either zero (IEEE 754 leaves it open)
                assert np.array_equal(got, want), (name, dt, j)
            else:
                assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), (name, dt, j)
            if oracle.have_ref():
                assert np.array_equal(oracle.map(name, [x, y], impl="ref").view(np.uint8), want.view(np.uint8))
