"""Functor plug-ins on the broadcast engine (b200phi_map) against the oracle and the golden
vectors made from the reference's own functors: compare / logical / bitwise over the
broadcast matrix of test_logical_op.py:45-60, the backward functors, the activation set."""
import numpy as np
import pytest
import torch

import cases
import gpu_util as G
import paddle2_b200 as pb
from paddle2_b200 import ops

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "phi_ref_golden.npz"))


@pytest.mark.parametrize("i", range(len(cases.MAP_SHAPES)))
def test_compare_logical_bitwise_golden(golden, i):
    _xs, _ys, src = cases.MAP_SHAPES[i]
    for dt in ("float32", "int32", "bool"):
        x, y = golden[f"map_{i}_{dt}_x"], golden[f"map_{i}_{dt}_y"]
        dx, dy = G.to_dev(x), G.to_dev(y)
        names = cases.COMPARE_OPS + cases.LOGICAL_OPS + ([] if dt == "float32" else cases.BITWISE_OPS)
        for name in names:
            want = golden[f"map_{i}_{dt}_{name}"]
            got = G.to_np(getattr(pb, name)(dx, dy))
            assert got.dtype == want.dtype and got.shape == want.shape, (src, name, dt)
            assert np.array_equal(got, want), (src, name, dt)


@pytest.mark.parametrize("dt", cases.MAP_ALL_DTYPES)
@pytest.mark.parametrize("shapes", [((257, 1031), (257, 1031)), ((64, 3, 500), (3, 1)),
                                    ((500, 77), (500, 1)), ((9, 1000), (1000,)), ((1, 4100), (33, 1))])
def test_compare_logical_every_dtype(oracle, dt, shapes):
    """aligned vector path, splat operand on either side, ragged tails; bit-exact."""
    rng = np.random.default_rng(31)
    x = cases.map_operand(rng, dt, shapes[0], "ties")
    y = cases.map_operand(rng, dt, shapes[1], "ties")
    dx, dy = G.to_dev(x), G.to_dev(y)
    names = cases.COMPARE_OPS + cases.LOGICAL_OPS
    if dt in cases.MAP_INT_DTYPES:
        names = names + cases.BITWISE_OPS
    for name in names:
        got = G.to_np(getattr(pb, name)(dx, dy))
        want = oracle.map(name, [x, y])
        assert got.dtype == want.dtype
        assert np.array_equal(got, want), (name, dt, shapes)
    for name in ["logical_not"] + (["bitwise_not"] if dt in cases.MAP_INT_DTYPES else []):
        assert np.array_equal(G.to_np(getattr(pb, name)(dx)), oracle.map(name, [x])), (name, dt)


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
def test_compare_16bit(oracle, tdt):
    """bf16/fp16 operands widen exactly, so every comparison is exact against the fp32 oracle."""
    rng = np.random.default_rng(5)
    x = torch.from_numpy((np.round(rng.random((65, 520)) * 8 - 4, 1)).astype(np.float32)).to(tdt)
    y = torch.from_numpy((np.round(rng.random((520,)) * 8 - 4, 1)).astype(np.float32)).to(tdt)
    xf, yf = x.float().numpy(), y.float().numpy()
    for name in cases.COMPARE_OPS + cases.LOGICAL_OPS:
        got = G.to_np(getattr(pb, name)(x.cuda(), y.cuda()))
        assert np.array_equal(got, oracle.map(name, [xf, yf])), name


def test_equal_tolerant_form():
    a = np.array([1e-9, 0.0, np.inf, np.inf, np.nan, 1.0, 1.0], dtype=np.float32)
    b = np.array([2e-9, 9e-9, np.inf, -np.inf, np.nan, 1.0, np.nextafter(np.float32(1), np.float32(2))],
                 dtype=np.float32)
    assert G.to_np(pb.equal(G.to_dev(a), G.to_dev(b))).tolist() == [True, True, True, False, False, True, False]
    assert G.to_np(pb.not_equal(G.to_dev(a), G.to_dev(b))).tolist() == [False, False, False, True, True, False, True]
    ad, bd = a.astype(np.float64), b.astype(np.float64)
    assert G.to_np(pb.equal(G.to_dev(ad), G.to_dev(bd))).tolist() == [True, True, True, False, False, True, False]


def test_backward_functors_golden(golden):
    x, y, g = golden["mapgrad_x"], golden["mapgrad_y"], golden["mapgrad_g"]
    for name in ("equal_mul", "max_grad_x", "max_grad_y", "min_grad_x", "min_grad_y", "div_grad_y"):
        got = G.to_np(ops._map(name, [G.to_dev(x), G.to_dev(y), G.to_dev(g)]))
        want = golden[f"mapgrad_{name}"]
        if name == "div_grad_y":
            np.testing.assert_allclose(got, want, rtol=1e-6)
        else:
            assert np.array_equal(got, want), name


@pytest.mark.parametrize("dt", ["float32", "float64", "int32", "int64"])
@pytest.mark.parametrize("shapes", [((64, 3000), (64, 1), (64, 1)),       # last-axis max_grad: x streamed
                                    ((300, 2048), (2048,), (2048,)),       # axis-0 max_grad: all streamed
                                    ((16, 33, 70), (16, 1, 70), (16, 33, 70)),
                                    ((7, 5, 3), (7, 1, 1), (1, 5, 1))])    # generic walk
def test_backward_functors_shapes(oracle, dt, shapes):
    rng = np.random.default_rng(77)
    x = cases.map_operand(rng, dt, shapes[0], "ties")
    y = cases.map_operand(rng, dt, shapes[1], "ties")
    g = cases.map_operand(rng, dt, shapes[2])
    names = ["equal_mul", "max_grad_x", "max_grad_y", "min_grad_x", "min_grad_y"]
    for name in names:
        got = G.to_np(ops._map(name, [G.to_dev(x), G.to_dev(y), G.to_dev(g)]))
        assert np.array_equal(got, oracle.map(name, [x, y, g])), (name, dt, shapes)
    got = G.to_np(ops._map("equal_t", [G.to_dev(x), G.to_dev(y)]))
    assert np.array_equal(got, oracle.map("equal_t", [x, y]))


@pytest.mark.parametrize("case", cases.ACTIVATIONS, ids=[c[0] for c in cases.ACTIVATIONS])
def test_activation_golden_and_shapes(oracle, golden, case):
    name, p0, p1, dom = case
    x = golden[f"act_{name}_x"]
    got = G.to_np(ops._map(name, [G.to_dev(x)], p0, p1))
    np.testing.assert_allclose(got, golden[f"act_{name}_y"], rtol=1e-5, atol=1e-6, err_msg=name)
    for shape in cases.ACT_SHAPES:
        for dt, rtol in (("float32", 1e-5), ("float64", 1e-12)):
            x = cases.map_operand(np.random.default_rng(9), dt, shape, dom)
            got = G.to_np(ops._map(name, [G.to_dev(x)], p0, p1))
            want = oracle.map(name, [x], p0, p1)
            assert got.shape == want.shape
            np.testing.assert_allclose(got, want, rtol=rtol, atol=rtol * 0.1, err_msg=f"{name} {dt} {shape}")


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("case", cases.ACTIVATIONS, ids=[c[0] for c in cases.ACTIVATIONS])
def test_activation_16bit(oracle, case, tdt):
    """16-bit storage, fp32 compute (MPType): rel 1e-2 against the fp32 oracle of the widened input."""
    name, p0, p1, dom = case
    x = torch.from_numpy(cases.map_operand(np.random.default_rng(11), "float32", [37, 264], dom)).to(tdt)
    got = ops._map(name, [x.cuda()], p0, p1)
    assert got.dtype == tdt
    want = oracle.map(name, [x.float().numpy()], p0, p1)
    np.testing.assert_allclose(got.float().cpu().numpy(), want, rtol=1e-2, atol=1e-2, err_msg=name)


def test_named_wrappers_and_attrs(oracle):
    x = cases.map_operand(np.random.default_rng(3), "float32", [19, 130], "wide")
    dx = G.to_dev(x)
    for fn, name, p0, p1 in [(lambda t: pb.leaky_relu(t, 0.1), "leaky_relu", 0.1, 0),
                             (lambda t: pb.elu(t, 0.7), "elu", 0.7, 0),
                             (lambda t: pb.softplus(t, 1.0, 20.0), "softplus", 1.0, 20.0),
                             (pb.relu6, "relu6", 6.0, 0), (pb.hardsigmoid, "hard_sigmoid", 0.1666667, 0.5),
                             (pb.swish, "swish", 1.0, 0), (pb.mish, "mish", 20.0, 0),
                             (pb.silu, "silu", 0, 0), (pb.abs, "abs", 0, 0)]:
        np.testing.assert_allclose(G.to_np(fn(dx)), oracle.map(name, [x], p0, p1), rtol=1e-5, atol=1e-6)
    xi = cases.map_operand(np.random.default_rng(3), "int64", [19, 130])
    assert np.array_equal(G.to_np(pb.abs(G.to_dev(xi))), np.abs(xi))


def test_errors_and_empty():
    from paddle2_b200._lib import B200PhiError
    for xs, ys, _src in cases.MAP_ERROR_SHAPES:
        with pytest.raises((B200PhiError, ValueError)):
            pb.logical_and(torch.zeros(xs, dtype=torch.bool).cuda(), torch.zeros(ys, dtype=torch.bool).cuda())
    with pytest.raises(B200PhiError):
        pb.bitwise_and(torch.zeros(3).cuda(), torch.zeros(3).cuda())
    with pytest.raises(B200PhiError):
        pb.exp(torch.zeros(3, dtype=torch.int32).cuda())
    assert pb.equal(torch.zeros((0, 4)).cuda(), torch.zeros((4,)).cuda()).shape == (0, 4)
    assert pb.exp(torch.zeros((0,)).cuda()).numel() == 0


def test_large_shape_properties():
    """Full-size checks without the oracle: exp/log round trip, x<y xor x>=y, De Morgan."""
    torch.manual_seed(0)
    x = torch.rand((4096, 8192), device="cuda") + 0.5
    y = torch.rand((8192,), device="cuda") + 0.5
    back = pb.exp(pb.log(x))
    assert torch.allclose(back, x, rtol=1e-5, atol=0)
    lt, ge = pb.less_than(x, y), pb.greater_equal(x, y)
    assert bool(torch.all(lt ^ ge))
    assert bool(torch.all(pb.logical_not(pb.logical_and(lt, ge)) == pb.logical_or(pb.logical_not(lt), pb.logical_not(ge))))
    assert int(lt.sum()) == int((x < y).sum())


@pytest.mark.parametrize("dt", cases.SHIFT_DTYPES)
def test_shifts_golden_and_shapes(oracle, golden, dt):
    """bitwise_left_shift / bitwise_right_shift, arithmetic and logical: counts below zero, in range
    and beyond the bit width (funcs/bitwise_functors.h:49-206); bit-exact."""
    x, y = golden[f"shift_{dt}_x"], golden[f"shift_{dt}_y"]
    for name in cases.SHIFT_OPS:
        got = G.to_np(ops._map(name, [G.to_dev(x), G.to_dev(y)]))
        assert np.array_equal(got, golden[f"shift_{dt}_{name}"]), (dt, name)
    rng = np.random.default_rng(2)
    bits = np.dtype(dt).itemsize * 8
    info = np.iinfo(dt)
    xx = rng.integers(info.min, info.max, size=(129, 1031), dtype=dt, endpoint=True)
    yy = rng.integers(0 if dt == "uint8" else -2, bits + 2, size=(1031,)).astype(dt)
    dx, dy = G.to_dev(xx), G.to_dev(yy)
    for arith in (True, False):
        assert np.array_equal(G.to_np(pb.bitwise_left_shift(dx, dy, arith)),
                              oracle.map("left_shift_arithmetic" if arith else "left_shift_logic", [xx, yy]))
        assert np.array_equal(G.to_np(pb.bitwise_right_shift(dx, dy, arith)),
                              oracle.map("right_shift_arithmetic" if arith else "right_shift_logic", [xx, yy]))


def test_unary_into_a_larger_shape():
    """a unary functor whose operand is broadcast (the expand form of reduce_grad.h:30-42): splat along
    the row, splat across rows, and a single element"""
    x = torch.linspace(-2, 2, 37)
    for shape, src in (((5, 37), x), ((37, 40), x.reshape(37, 1)), ((3, 7, 9), x[:1].reshape(1, 1, 1)),
                       ((6, 1, 37), x.reshape(1, 1, 37))):
        got = ops._map("exp", [src.cuda()], out_shape=list(shape)).cpu()
        assert torch.allclose(got, torch.exp(src).expand(shape), rtol=1e-6)
        gotb = ops._map("logical_not", [src.cuda()], out_shape=list(shape)).cpu()
        assert torch.equal(gotb, (src == 0).expand(shape))


@pytest.mark.parametrize("name,dtypes", cases.ARITH_OPS)
def test_arith_binary_functors(oracle, golden, name, dtypes):
    """The arithmetic binaries the reference registers next to add .. minimum
    (kps/elementwise_kernel.cu:183-293): golden arrays from its own functors.  Integer results,
    copysign / fmax / fmin / heaviside / remainder / floor_divide are bit-exact; pow within 2 ulp
    (device pow vs libm)."""
    for dt in dtypes:
        for j in range(len(cases.ARITH_SHAPES)):
            x, y = golden[f"arith_{name}_{dt}_{j}_x"], golden[f"arith_{name}_{dt}_{j}_y"]
            want = golden[f"arith_{name}_{dt}_{j}_out"]
            got = G.to_np(getattr(pb, name)(G.to_dev(x), G.to_dev(y)))
            assert got.dtype == want.dtype and got.shape == want.shape, (name, dt, j)
            if name == "elementwise_pow" and dt.startswith("float"):
                np.testing.assert_allclose(got, want, rtol=4e-7 if dt == "float32" else 1e-15)
            else:
                assert np.array_equal(got, want, equal_nan=True), (name, dt, j)
                if dt.startswith("float") and name not in ("fmax", "fmin"):   # signed zeros too (fmax(+0,-0) is open)
                    assert np.array_equal(np.signbit(got), np.signbit(want)), (name, dt, j)


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("name", ["remainder", "floor_divide", "elementwise_pow", "copysign", "fmax", "fmin", "heaviside"])
def test_arith_binary_16bit_and_large(oracle, name, tdt):
    """16-bit floats compute in fp32 and round once (the reference's float16 / bfloat16
    specialisations); a large aligned shape and a channel broadcast exercise the vector paths."""
    rng = np.random.default_rng(17)
    for xs, ys in (((257, 1032), (257, 1032)), ((64, 8, 196), (8, 1))):
        x32, y32 = cases.arith_operands(rng, name, "float32", xs, ys)
        xt, yt = torch.from_numpy(x32).cuda().to(tdt), torch.from_numpy(y32).cuda().to(tdt)
        xw, yw = xt.float().cpu().numpy(), yt.float().cpu().numpy()
        if name in ("remainder", "floor_divide"):
            yw[yw == 0] = 1.5
            yt = torch.from_numpy(yw).cuda().to(tdt)
        want = torch.from_numpy(oracle.map(name, [xw, yw])).to(tdt)
        got = getattr(pb, name)(xt, yt).cpu()
        if name == "elementwise_pow":
            torch.testing.assert_close(got.float(), want.float(), rtol=1e-2, atol=0)
        else:
            assert torch.equal(got, want), (name, tdt, xs)
