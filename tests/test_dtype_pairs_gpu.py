"""The dtype combinations on the reference's registration lists that round 1 refused
(kps/reduce_kernel.cu:345-397, gpu/reduce.h:48-69): out_dtype != x.dtype for sum into 16-bit floats,
max / min / prod with another output dtype, complex64/128 sum / mean / add / subtract, float8 max.

The reference's rule for out_dtype != x.dtype is "cast x to the output dtype, then reduce"
(phi::Cast pre-pass); the oracle restates exactly that on the CPU."""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_util as G
import paddle2_b200 as pb
from paddle2_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _kind(x, axes, op, out_dtype):
    d = L.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = op, pb.ops._phi_dtype(x.dtype), pb.ops._phi_dtype(out_dtype), x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.naxes = len(axes)
    for i, a in enumerate(axes):
        d.axes[i] = a
    d.x = x.data_ptr()
    return L.lib().b200phi_reduce_plan_kind(C.byref(d))


@pytest.mark.parametrize("shape,axes", [((300, 257), [1]), ((4000, 96), [0]), ((12, 50, 33), [0, 2]), ((70000,), [0])])
@pytest.mark.parametrize("out", [torch.bfloat16, torch.float16])
def test_sum_into_16bit_from_fp32_rounds_each_element_first(shape, axes, out):
    """fp32 -> bf16 / fp16 sum: every element is rounded to the 16-bit type BEFORE it is added (the
    Cast pre-pass), accumulated in fp32, and the result rounded once.  Checked against torch doing
    exactly that in float64 accumulation (the fp32 accumulation error is far below the 16-bit ulp)."""
    g = torch.Generator(device="cuda").manual_seed(sum(shape))
    x = torch.rand(shape, device="cuda", generator=g) * 4 - 1
    assert _kind(x, axes, L.SUM, out) == 12          # kCastFirst
    got = pb.sum(x, axis=axes, dtype=out)
    assert got.dtype == out
    want = x.to(out).double().sum(dim=axes).to(out)
    tol = G.BF16_RTOL if out == torch.bfloat16 else 2e-3
    torch.testing.assert_close(got.float(), want.float(), rtol=tol, atol=tol)
    # and NOT the sum of the unrounded values when the two differ visibly
    exact = x.double().sum(dim=axes)
    rounded = x.to(out).double().sum(dim=axes)
    far = (exact - rounded).abs() > 4 * (rounded.abs() * (2.0 ** -8 if out == torch.bfloat16 else 2.0 ** -11))
    if bool(far.any()):
        assert bool(((got.double() - rounded).abs()[far] <= (got.double() - exact).abs()[far]).all())
    # mean follows the same rule
    gm = pb.ops._reduce(L.MEAN, x, axes, False, out)
    wm = (x.to(out).double().mean(dim=axes)).to(out)
    torch.testing.assert_close(gm.float(), wm.float(), rtol=tol, atol=tol)


@pytest.mark.parametrize("shape,axes", [((300, 257), [1]), ((4000, 96), [0]), ((12, 50, 33), [0, 2])])
def test_max_min_with_another_output_dtype(oracle, shape, axes):
    rng = np.random.default_rng(sum(shape))
    xf = (rng.standard_normal(shape) * 100).astype(np.float32)
    xt = G.to_dev(xf)
    for out, npd in ((torch.float64, np.float64), (torch.bfloat16, None), (torch.float16, np.float16)):
        got = pb.max(xt, axis=axes, dtype=out)
        assert got.dtype == out
        want = torch.from_numpy(oracle.reduce("max", xf, axes)).to(out)      # rounding is monotonic
        assert torch.equal(got.cpu(), want), out
        assert torch.equal(pb.min(xt, axis=axes, dtype=out).cpu(), torch.from_numpy(oracle.reduce("min", xf, axes)).to(out))
    xd = rng.standard_normal(shape)
    assert torch.equal(pb.max(G.to_dev(xd), axis=axes, dtype=torch.float32).cpu(),
                       torch.from_numpy(oracle.reduce("max", xd, axes)).float())
    xi = rng.integers(-(2 ** 30), 2 ** 30, size=shape).astype(np.int32)
    got = pb.max(G.to_dev(xi), axis=axes, dtype=torch.int64)
    assert got.dtype == torch.int64 and np.array_equal(G.to_np(got), oracle.reduce("max", xi, axes).astype(np.int64))
    # narrowing integers wrap element by element first (the cast), then reduce
    xl = rng.integers(-(2 ** 40), 2 ** 40, size=shape).astype(np.int64)
    got = pb.max(G.to_dev(xl), axis=axes, dtype=torch.int32)
    assert np.array_equal(G.to_np(got), oracle.reduce("max", xl.astype(np.int32), axes))
    got = pb.sum(G.to_dev(xl), axis=axes, dtype=torch.int32)
    assert np.array_equal(G.to_np(got), oracle.reduce("sum", xl.astype(np.int32), axes))


def test_prod_with_another_output_dtype(oracle):
    rng = np.random.default_rng(3)
    x = (0.9 + 0.2 * rng.random((40, 7, 30))).astype(np.float32)
    xt = G.to_dev(x)
    d = L.ReduceDesc()
    out = torch.empty(40, 30, device="cuda", dtype=torch.float64)
    d.op, d.in_dtype, d.out_dtype, d.rank = L.PROD, L.FLOAT32, L.FLOAT64, 3
    d.dims[0], d.dims[1], d.dims[2] = 40, 7, 30
    d.naxes = 1
    d.axes[0] = 1
    d.x, d.out = xt.data_ptr(), out.data_ptr()
    lib = L.lib()
    assert lib.b200phi_reduce_plan_kind(C.byref(d)) == 12
    ws_bytes = lib.b200phi_reduce_workspace(C.byref(d))
    assert ws_bytes >= x.size * 8
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws_bytes
    L.check(lib.b200phi_reduce(C.byref(d), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    np.testing.assert_allclose(G.to_np(out), oracle.reduce("prod", x.astype(np.float64), [1]), rtol=1e-12)


@pytest.mark.parametrize("dt", [torch.complex64, torch.complex128])
@pytest.mark.parametrize("shape,axes", [((5, 6, 10), [0]), ((3, 4), [0]), ((300, 257), [1]), ((4000, 96), [0]),
                                        ((12, 50, 33), [0, 2]), ((70000,), None), ((), None), ((7, 1, 9), [1])])
def test_complex_sum_and_mean(dt, shape, axes):
    """test_reduce_op.py:76-79 (complex128 (3,4) over [0]) and larger: component-wise sums."""
    g = torch.Generator(device="cuda").manual_seed(len(shape) + 1)
    rdt = torch.float32 if dt == torch.complex64 else torch.float64
    x = torch.complex(torch.rand(shape, device="cuda", generator=g, dtype=rdt) - 0.5,
                      torch.rand(shape, device="cuda", generator=g, dtype=rdt) - 0.5)
    got = pb.sum(x, axis=axes)
    assert got.dtype == dt
    want = x.to(torch.complex128).sum(dim=axes) if axes is not None else x.to(torch.complex128).sum()
    scale = x.abs().double().sum(dim=axes) if axes is not None else x.abs().double().sum()
    tol = 1e-5 if dt == torch.complex64 else 1e-12
    assert bool(((got.to(torch.complex128) - want).abs() <= tol * scale + 1e-30).all())
    gm = pb.mean(x, axis=axes)
    wm = x.to(torch.complex128).mean(dim=axes) if axes is not None else x.to(torch.complex128).mean()
    assert bool(((gm.to(torch.complex128) - wm).abs() <= tol * (wm.abs() + 1)).all())
    assert tuple(pb.sum(x, axis=axes, keepdim=True).shape) == tuple(
        x.sum(dim=axes, keepdim=True).shape if axes is not None else [1] * x.dim())


@pytest.mark.parametrize("dt", [torch.complex64, torch.complex128])
def test_complex_add_subtract_with_broadcast(dt):
    """test_elementwise_add_op.py:744-770: complex operands, numpy broadcasting and the legacy axis."""
    g = torch.Generator(device="cuda").manual_seed(4)
    rdt = torch.float32 if dt == torch.complex64 else torch.float64

    def cx(*shape):
        return torch.complex(torch.rand(shape, device="cuda", generator=g, dtype=rdt),
                             torch.rand(shape, device="cuda", generator=g, dtype=rdt))

    for xs, ys in (((13, 17), (13, 17)), ((2, 3, 100), (100,)), ((2, 100, 3), (100, 1)), ((10, 3, 1, 4), (10, 1, 12, 1))):
        x, y = cx(*xs), cx(*ys)
        assert torch.equal(pb.add(x, y), x + y)
        assert torch.equal(pb.subtract(x, y), x - y)
    x, y = cx(2, 100, 3), cx(100)
    assert torch.equal(pb.add(x, y, axis=1), x + y[None, :, None])
    x, y = cx(100, 2, 3), cx(100)
    assert torch.equal(pb.subtract(x, y, axis=0), x - y[:, None, None])
    with pytest.raises(pb.B200PhiError):
        pb.multiply(x, x)


@pytest.mark.parametrize("dt", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("shape,axes", [((300, 257), [1]), ((5000, 96), [0]), ((12, 50, 33), [0, 2]),
                                        ((1 << 22,), None), ((64, 4096), [1])])
def test_float8_max(dt, shape, axes):
    g = torch.Generator(device="cuda").manual_seed(9)
    x = (torch.randn(shape, device="cuda", generator=g) * 8).to(dt)
    got = pb.max(x, axis=axes)
    assert got.dtype == dt
    want = x.float().amax(dim=axes) if axes is not None else x.float().max()
    assert torch.equal(got.float(), want)          # the maximum of representable values is representable
    with pytest.raises(pb.B200PhiError):
        pb.sum(x, axis=axes)
