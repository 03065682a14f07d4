"""GPU parity of the reduction family against the oracle, through the C ABI
(paddle2_b200.ops -> ctypes -> libb200phi.so).  Replays the reference's own test
shapes (tests/cases.py) plus kernel-selection edge cases."""
import numpy as np
import pytest
import torch

import cases
import gpu_util as G
import paddle2_b200 as pb

pytestmark = pytest.mark.gpu

RT = {np.dtype(np.float32): G.FP32_RTOL, np.dtype(np.float64): G.FP64_RTOL}


def _run(op, xt, axis, keepdim):
    return getattr(pb, op)(xt, axis=axis, keepdim=keepdim)


@pytest.mark.parametrize("i", range(len(cases.REDUCE_CASES)))
def test_replay_reference_cases(oracle, i):
    op, dt, shape, axis, keepdim, src = cases.REDUCE_CASES[i]
    rng = np.random.default_rng(123 + i)
    kind = "logical" if op in ("any", "all") else ("prod" if op == "prod" else "uniform")
    x = cases.make_input(rng, dt, shape, kind)
    want = oracle.reduce(op, x, axis, keepdim)
    got = G.to_np(_run(op, G.to_dev(x), axis, keepdim))
    assert got.dtype == want.dtype, src
    G.assert_reduce_close(got, want, x, axis, op, RT.get(np.dtype(dt), 0))


@pytest.mark.parametrize("i", range(len(cases.SUM_DTYPE_CASES)))
def test_sum_with_dtype(oracle, i):
    idt, odt, shape, axis, keepdim, src = cases.SUM_DTYPE_CASES[i]
    x = cases.make_input(np.random.default_rng(1000 + i), idt, shape)
    want = oracle.reduce("sum", x, axis, keepdim, out_dtype=odt)
    got = G.to_np(pb.sum(G.to_dev(x), axis=axis, dtype=odt, keepdim=keepdim))
    assert got.dtype == want.dtype, src
    G.assert_reduce_close(got, want, x, axis, "sum", RT.get(np.dtype(odt), 0))


def test_sum_default_promotion(oracle):
    # int32 / bool sum to int64 by default (test_reduce_op.py:1731)
    x = np.random.default_rng(5).integers(-100, 100, size=(2, 3, 4)).astype(np.int32)
    got = pb.sum(G.to_dev(x))
    assert got.dtype == torch.int64 and int(got) == int(x.astype(np.int64).sum())
    b = np.random.default_rng(6).integers(0, 2, size=(33, 77)).astype(np.bool_)
    got = pb.sum(G.to_dev(b), axis=0)
    assert got.dtype == torch.int64
    assert np.array_equal(G.to_np(got), oracle.reduce("sum", b, [0], out_dtype=np.int64))


# every kernel kind and its boundaries: (shape, axis)
KIND_SHAPES = [
    ((3000, 10), [1]),          # short rows: 2 lanes per row
    ((100000, 3), [1]),         # short rows: 1 lane per row
    ((5000, 100), [1]),         # short rows: 16 lanes per row
    ((70, 128), [1]),           # short rows, partial last warp
    ((40, 5), [1]),             # fewer than 64 rows: warp per row
    ((2400, 1027), [1]),        # warp-per-row, misaligned rows (head/tail peel)
    ((4096, 4096), [1]),        # warp-per-row, 16 KiB rows (C1 shape class)
    ((512, 20000), [1]),        # CTA-per-row
    ((300, 8193), [1]),         # CTA-per-row, odd length
    ((3, 1 << 20), [1]),        # few long rows -> split across CTAs + last-block finalise
    ((1 << 22,), None),         # full reduction -> split
    ((5000, 256), [0]),         # column reduce, vector path, row split
    ((4099, 33), [0]),          # column reduce, scalar path (inner not a vector multiple)
    ((7, 300, 64), [1]),        # [outer, R, inner]
    ((64, 40, 3), [1]),         # tiny inner
    ((6, 5, 40, 7), [1, 3]),    # generic: reduce dims interleaved, inner reduced
    ((6, 50, 4, 70), [0, 2]),   # generic: inner kept
    ((100000, 4), [0]),         # one vector per row, everything over ty
]


@pytest.mark.parametrize("shape,axis", KIND_SHAPES)
@pytest.mark.parametrize("dt", ["float32", "int32", "int64", "float64"])
def test_kernel_kinds(oracle, shape, axis, dt):
    rng = np.random.default_rng(abs(hash((shape, dt))) % (2 ** 31))
    if np.issubdtype(np.dtype(dt), np.integer):
        x = rng.integers(-1000, 1000, size=shape).astype(dt)
    else:
        x = rng.random(shape).astype(dt)
    for op in ("sum", "max", "min", "mean"):
        if op == "mean" and np.issubdtype(np.dtype(dt), np.integer):
            continue
        out_dt = oracle.sum_out_dtype(dt) if op == "sum" else np.dtype(dt)
        want = oracle.reduce(op, x, axis, False, out_dtype=out_dt)
        got = G.to_np(_run(op, G.to_dev(x), axis, False))
        assert got.dtype == want.dtype
        G.assert_reduce_close(got, want, x, axis, op, RT.get(np.dtype(dt), 0))


@pytest.mark.parametrize("shape,axis", [((2400, 1027), [1]), ((512, 20000), [1]), ((5000, 256), [0]),
                                        ((4099, 33), [0]), ((6, 5, 40, 7), [1, 3]), ((1 << 20,), None)])
def test_bf16_fp16_fp32_accumulate(oracle, shape, axis):
    rng = np.random.default_rng(11)
    x32 = G.bf16_round(rng.random(shape).astype(np.float32))
    xt = torch.from_numpy(x32).cuda().bfloat16()
    for op in ("sum", "mean", "max", "min"):
        want = oracle.reduce(op, x32, axis, False)       # fp32 oracle on the widened input
        got = _run(op, xt, axis, False)
        assert got.dtype == torch.bfloat16
        got = G.to_np(got)
        if op in ("max", "min"):
            assert np.array_equal(got, want)             # max of representable values: exact
        else:
            np.testing.assert_allclose(got, want, rtol=G.BF16_RTOL)
    # bf16 -> fp32 output keeps the fp32 accumulator (kps/reduce_kernel.cu:251-265)
    got = G.to_np(pb.sum(xt, axis=axis, dtype="float32"))
    want = oracle.reduce("sum", x32, axis, False)
    G.assert_reduce_close(got, want, x32, axis, "sum", G.FP32_RTOL)
    xh = torch.from_numpy(x32).cuda().half()   # mean, not sum: 2^20 halves overflow fp16's 65504
    got = G.to_np(pb.mean(xh, axis=axis).float())
    np.testing.assert_allclose(got, oracle.reduce("mean", xh.float().cpu().numpy(), axis, False),
                               rtol=G.FP16_RTOL)


def test_small_int_and_bool_types(oracle):
    rng = np.random.default_rng(3)
    for dt in ("uint8", "int8", "int16"):
        x = rng.integers(0, 100, size=(77, 513)).astype(dt)
        for axis in ([0], [1], None):
            want = oracle.reduce("sum", x, axis, False, out_dtype=np.int64)
            got = G.to_np(pb.sum(G.to_dev(x), axis=axis, dtype="int64"))
            assert np.array_equal(got, want)
            assert np.array_equal(G.to_np(pb.max(G.to_dev(x), axis=axis)), oracle.reduce("max", x, axis))
    b = rng.integers(0, 2, size=(33, 1000)).astype(np.bool_)
    b[5] = True
    b[7] = False
    for axis in ([0], [1], None):
        for op in ("any", "all"):
            assert np.array_equal(G.to_np(_run(op, G.to_dev(b), axis, False)), oracle.reduce(op, b, axis))


def test_determinism_and_workspace_reuse():
    """Last-block finalise: same bits on every run, tickets reset for the next call."""
    x = torch.rand(1 << 22, device="cuda")
    a = pb.sum(x)
    for _ in range(5):
        assert torch.equal(pb.sum(x), a)
    y = torch.rand(20000, 512, device="cuda")
    c = pb.sum(y, axis=0)
    for _ in range(5):
        assert torch.equal(pb.sum(y, axis=0), c)


def test_errors_are_loud():
    with pytest.raises(pb.B200PhiError, match="must not empty"):
        pb.sum(torch.empty(0, 3, device="cuda"))
    with pytest.raises(ValueError):
        pb.sum(torch.ones(2, 3, device="cuda"), axis=5)
    with pytest.raises(TypeError):
        pb.sum(torch.ones(2, 3, device="cuda").to(torch.uint16))   # no phi kernel for uint16 either


def test_cast_shortcut(oracle):
    x = np.random.default_rng(2).random((1000, 37)).astype(np.float32) * 100
    assert np.array_equal(G.to_np(pb.cast(G.to_dev(x), "int64")), x.astype(np.int64))
    assert np.array_equal(G.to_np(pb.cast(G.to_dev(x), "float64")), x.astype(np.float64))


# ---- per-channel pattern [A, K, B] with A and B reduced (two-stage plan) -------------------------
@pytest.mark.parametrize("shape,axis", [((16, 8, 14, 14), [0, 2, 3]), ((33, 5, 1000), [0, 2]), ((2, 64, 4099), (0, 2)),
                                        ((300, 3, 17), [0, 2]), ((4, 6, 1, 32), [0, 3])])
@pytest.mark.parametrize("dt", ["float32", "float64", "int32", "int64", "bool"])
def test_per_channel_reductions(oracle, shape, axis, dt):
    """sum / mean / max / min / prod / any / all over (0, 2, ...) keeping the channel axis: the row
    kernel into scratch, then the column kernel.  Integer and bool results are bit-exact."""
    rng = np.random.default_rng(11)
    axis = list(axis)
    ops_for = {"float32": ["sum", "mean", "max", "min", "prod", "any", "all"],
               "float64": ["sum", "mean", "max", "prod"], "int32": ["sum", "max", "min", "any"],
               "int64": ["sum", "min", "prod", "all"], "bool": ["sum", "any", "all"]}[dt]
    for op in ops_for:
        kind = "logical" if op in ("any", "all") else ("prod" if op == "prod" else "uniform")
        x = cases.make_input(rng, dt, shape, kind)
        if op == "prod" and dt.startswith("float"):
            x = (0.9 + 0.2 * rng.random(shape)).astype(dt)
        for keepdim in (False, True):
            odt = oracle.sum_out_dtype(x.dtype) if op == "sum" else None   # int32 / bool sum into int64
            want = oracle.reduce(op, x, axis, keepdim, out_dtype=odt)
            got = G.to_np(getattr(pb, op)(G.to_dev(x), axis=axis, keepdim=keepdim))
            assert got.shape == want.shape and got.dtype == want.dtype, (op, dt)
            if dt.startswith("float") and op == "prod":
                # n factors = n roundings in either order: the two products drift apart like sqrt(n) ulp
                n = x.size // max(want.size, 1)
                eps = 6e-8 if dt == "float32" else 1.2e-16
                np.testing.assert_allclose(got, want, rtol=max(G.FP32_RTOL if dt == "float32" else 1e-12,
                                                               4 * eps * np.sqrt(n)))
            elif dt.startswith("float") and op in ("sum", "mean"):
                G.assert_reduce_close(got, want, x, axis, op, G.FP32_RTOL if dt == "float32" else 1e-12)
            else:
                assert np.array_equal(got, want), (op, dt, shape)


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
def test_per_channel_16bit(tdt):
    x = torch.randn((64, 48, 28, 28), generator=torch.Generator().manual_seed(3)).to(tdt)
    x64 = x.double()
    got = pb.mean(x.cuda(), axis=[0, 2, 3])
    assert got.dtype == tdt and tuple(got.shape) == (48,)
    np.testing.assert_allclose(got.double().cpu().numpy(), x64.mean(dim=(0, 2, 3)).numpy(), rtol=1e-2, atol=1e-3)
    got = pb.sum(x.cuda(), axis=[0, 2, 3], dtype="float32")      # bf16 in, fp32 out: the tighter check
    np.testing.assert_allclose(got.cpu().numpy(), x64.sum(dim=(0, 2, 3)).numpy(), rtol=1e-4, atol=1e-2)
    assert torch.equal(pb.max(x.cuda(), axis=[0, 2, 3]).cpu(), x.amax(dim=(0, 2, 3)))
    assert torch.equal(pb.min(x.cuda(), axis=[0, 2, 3]).cpu(), x.amin(dim=(0, 2, 3)))


def test_per_channel_full_size_properties():
    """[256,64,112,112] fp32 over (0,2,3): equals the two one-axis reductions it stands for."""
    x = torch.rand((256, 64, 112, 112), device="cuda")
    got = pb.sum(x, axis=[0, 2, 3])
    ref = pb.sum(pb.sum(x.reshape(256 * 64, 112 * 112), axis=-1).reshape(256, 64), axis=0)
    assert torch.allclose(got, ref, rtol=1e-6)
    assert torch.allclose(got.double(), x.double().sum(dim=(0, 2, 3)), rtol=1e-5)
    assert torch.equal(pb.max(x, axis=[0, 2, 3]), x.amax(dim=(0, 2, 3)))
