"""Reductions over general keep/reduce interleavings (the reference's ReduceAnyKernel +
IndexCalculator, reduce_function.h:573-699): tile_keep_kernel / tile_red_kernel against the
oracle, on the shapes the reference's own tests pin (test_reduce_op.py:464-484, 1262-1264) and on
shapes large enough to split the reduce space across CTAs.  Every case first asserts that the
shape really takes a tiled plan, so a silent fall-back to the scalar index walk would fail."""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_util as G
import paddle2_b200 as pb
from paddle2_b200 import _lib as L

pytestmark = pytest.mark.gpu

KEEP, RED, GENERIC = 10, 11, 4


def plan_kind(x, axes, op=L.SUM, out_dtype=0):
    d = L.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = op, pb.ops._phi_dtype(x.dtype), out_dtype, x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.naxes = len(axes)
    for i, a in enumerate(axes):
        d.axes[i] = a
    d.x = x.data_ptr()
    return L.lib().b200phi_reduce_plan_kind(C.byref(d)), L.lib().b200phi_reduce_workspace(C.byref(d))


# (shape, axes, expected kind): keep = innermost merged dim kept, red = innermost reduced
PATTERNS = [
    ((2, 4, 3, 5, 6, 10), [2, 4], KEEP),                    # test_reduce_op.py:464-466
    ((2, 4, 3, 5, 6, 3, 2, 4), [2, 3, 4], None),            # :482-484 (merges to [K,R,K]: column kind)
    ((2, 5, 3, 2, 2, 3, 4, 2), [4, 2, 3], None),            # :1262-1264
    ((1, 3, 1, 2, 1, 4, 3, 10), [0, 3], KEEP),              # :123-126  -> [K3,R2,K120]: R < 16
    ((16, 24, 56, 56), [0, 2], None),                       # [R,K,R,K]: peels the 56-row dim, then a column reduce
    ((16, 24, 56, 56), [1, 3], None),                       # [K,R,K,R]: peels the contiguous run, then ...
    ((12, 24, 10, 14), [0, 2], KEEP),                       # [R,K,R,K] with short dims: single-pass tiled
    ((12, 24, 10, 14), [1, 3], None),                       # (a 24-long fp32 dim is still worth peeling)
    ((12, 8, 10, 12), [1, 3], RED),                         # [K,R,K,R] with short dims
    ((6, 5, 7, 9, 11, 12), [0, 2, 4], KEEP),                # [R,K,R,K,R,K]
    ((6, 5, 7, 9, 11, 12), [1, 3, 5], RED),                 # [K,R,K,R,K,R]
    ((3, 4, 5, 3, 4, 5, 3, 4, 8), [0, 2, 4, 6, 8], RED),    # rank 9, five reduced dims
    ((3, 4, 5, 3, 4, 5, 3, 4, 8), [1, 3, 5, 7], KEEP),      # rank 9, five kept dims
    ((3000, 7, 12), [0, 2], None),                          # [R,K,R] short inner run: peels the 3000-row dim
    ((12, 7000, 12), [0, 2], RED),                          # [R,K,R], both reduced dims short
    ((9, 40000, 20), [0], KEEP),                            # [R,K] with R < 16
    ((50, 9, 4000), [1], KEEP),                             # [K,R,K] with R < 16
    ((7, 33, 5, 129), [0, 2], KEEP),                        # odd inner extent: scalar walk of the kept dim
    ((7, 33, 5, 129), [1, 3], None),                        # odd run length 129: long enough to peel (row kernel)
    ((7, 13, 5, 15), [1, 3], RED),                          # odd run length, short: scalar walk of the run
    ((40, 3, 70000), [0, 2], None),                         # [R,K,R] long run: two-stage
]


@pytest.mark.parametrize("shape,axes,kind", PATTERNS)
@pytest.mark.parametrize("op", ["sum", "mean", "max", "min", "prod"])
def test_patterns_fp32_against_the_oracle(oracle, shape, axes, kind, op):
    rng = np.random.default_rng(abs(hash((shape, tuple(axes)))) % (2 ** 31))
    x = rng.random(shape).astype(np.float32)
    if op == "prod":
        x = (0.9 + 0.2 * x).astype(np.float32)
    xt = G.to_dev(x)
    k, _ws = plan_kind(xt, axes, getattr(L, op.upper()))
    if kind is not None:
        assert k == kind, f"expected plan kind {kind}, got {k}"
    assert k != GENERIC, "no shape of this list may fall back to the scalar index walk"
    got = G.to_np(getattr(pb, op)(xt, axis=axes))
    want = oracle.reduce(op, x, axes)
    rtol = G.FP32_RTOL
    if op == "prod":  # n rounded multiplications in either order: the bound grows with n
        rtol = max(rtol, float(np.prod([shape[a] for a in axes])) * 2.0 ** -24)
    G.assert_reduce_close(got, want, x, axes, op, rtol)
    got_k = G.to_np(getattr(pb, op)(xt, axis=axes, keepdim=True)) if op != "prod" else None
    if got_k is not None:
        assert np.array_equal(got_k.reshape(got.shape), got)


@pytest.mark.parametrize("shape,axes", [((16, 24, 56, 56), [0, 2]), ((16, 24, 56, 56), [1, 3]),
                                        ((6, 5, 7, 9, 11, 12), [0, 2, 4]), ((7, 33, 5, 129), [1, 3]),
                                        ((3000, 7, 12), [0, 2]), ((9, 40000, 20), [0])])
def test_patterns_other_dtypes(oracle, shape, axes):
    rng = np.random.default_rng(sum(shape))
    # integers: bit-exact; int32 sum promotes to int64 (SumRawInferMeta)
    xi = rng.integers(-50, 50, size=shape).astype(np.int32)
    assert np.array_equal(G.to_np(pb.sum(G.to_dev(xi), axis=axes)), oracle.reduce("sum", xi, axes, out_dtype=np.int64))
    assert np.array_equal(G.to_np(pb.max(G.to_dev(xi), axis=axes)), oracle.reduce("max", xi, axes))
    xl = rng.integers(-(2 ** 40), 2 ** 40, size=shape).astype(np.int64)
    assert np.array_equal(G.to_np(pb.min(G.to_dev(xl), axis=axes)), oracle.reduce("min", xl, axes))
    # bool any / all
    xb = rng.random(shape) > 0.97
    assert np.array_equal(G.to_np(pb.any(G.to_dev(xb), axis=axes)), oracle.reduce("any", xb, axes))
    assert np.array_equal(G.to_np(pb.all(G.to_dev(~xb), axis=axes)), oracle.reduce("all", ~xb, axes))
    # fp64
    xd = rng.random(shape)
    np.testing.assert_allclose(G.to_np(pb.sum(G.to_dev(xd), axis=axes)), oracle.reduce("sum", xd, axes), rtol=1e-12)
    # bf16 / fp16: widened oracle, rel 1e-2; max exact
    x32 = G.bf16_round(rng.random(shape).astype(np.float32))
    xt = torch.from_numpy(x32).cuda().bfloat16()
    k, _ = plan_kind(xt, axes)
    assert k != GENERIC
    np.testing.assert_allclose(G.to_np(pb.sum(xt, axis=axes)), oracle.reduce("sum", x32, axes), rtol=G.BF16_RTOL)
    assert np.array_equal(G.to_np(pb.max(xt, axis=axes)), oracle.reduce("max", x32, axes))
    xh = torch.from_numpy(x32).cuda().half()
    np.testing.assert_allclose(G.to_np(pb.mean(xh, axis=axes)).astype(np.float32),
                               oracle.reduce("mean", xh.float().cpu().numpy(), axes), rtol=G.FP16_RTOL)
    # norm-type reductions ride the same kernels (TransformOp slot)
    xf = (rng.random(shape).astype(np.float32) - 0.5)
    if len(axes) == 1:
        np.testing.assert_allclose(G.to_np(pb.ops.p_norm(G.to_dev(xf), 2.0, axes[0])),
                                   oracle.p_norm(xf, 2.0, axes[0]), rtol=G.FP32_RTOL * 10)


@pytest.mark.parametrize("shape,axes,kind", [((64, 64, 112, 112), [0, 2], None), ((64, 64, 112, 112), [1, 3], None),
                                             ((8, 64, 48, 80, 6, 10), [2, 4], None),
                                             ((4, 5, 6, 7, 8, 9, 10, 11), [2, 3, 4], None),
                                             ((8, 300000, 3, 8), [0, 3], RED), ((12, 3000000), [0], KEEP),
                                             ((12, 40, 12, 50000), [0, 2], KEEP), ((1500, 12, 900, 12), [1, 3], RED)])
def test_large_shapes_split_the_reduce_space(shape, axes, kind):
    """Big enough that the reduced extent is cut across CTAs (workspace > 0 where outputs are few):
    fp32 sums against float64 torch, max bit-exact, determinism across two calls."""
    g = torch.Generator(device="cuda").manual_seed(len(shape) * 7 + axes[0])
    x = torch.rand(shape, device="cuda", generator=g)
    k, ws = plan_kind(x, axes)
    if kind is not None:
        assert k == kind
    assert k != GENERIC
    got = pb.sum(x, axis=axes)
    want = x.double().sum(dim=axes)
    scale = x.double().abs().sum(dim=axes)
    assert bool(((got.double() - want).abs() <= G.FP32_RTOL * scale).all())
    assert torch.equal(pb.sum(x, axis=axes), got)                       # deterministic finalise
    assert torch.equal(pb.max(x, axis=axes), x.amax(dim=axes))
    assert torch.equal(pb.min(x, axis=axes), x.amin(dim=axes))
    xb = x.bfloat16()
    gb = pb.sum(xb, axis=axes, dtype="float32")
    wb = xb.double().sum(dim=axes)
    assert bool(((gb.double() - wb).abs() <= G.FP32_RTOL * xb.double().abs().sum(dim=axes)).all())


def test_unaligned_base_pointer_takes_the_scalar_walk():
    flat = torch.rand(16 * 24 * 56 * 56 + 1, device="cuda")
    x = flat[1:].view(16, 24, 56, 56)
    for axes in ([0, 2], [1, 3]):
        assert plan_kind(x, axes)[0] != GENERIC
        got = pb.sum(x, axis=axes)
        want = x.double().sum(dim=axes)
        assert bool(((got.double() - want).abs() <= G.FP32_RTOL * want.abs()).all())


def test_tiled_kernels_on_the_peelable_shapes_too():
    """B200PHI_NO_TWO_STAGE=1 (read once per process) keeps every pattern on the single-pass tiled
    kernels, so they are also checked on the long-dim shapes that normally take two stages."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = f"""
import sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {os.path.join(root, 'tests')!r})
import torch, paddle2_b200 as pb
import test_tiled_reduce_gpu as T
g = torch.Generator(device='cuda').manual_seed(2)
for shape, axes, kind in [((64, 64, 112, 112), [0, 2], T.KEEP), ((64, 64, 112, 112), [1, 3], T.RED),
                          ((8, 64, 48, 80, 6, 10), [2, 4], T.KEEP), ((6, 5, 7, 9, 11, 12), [1, 3, 5], T.RED),
                          ((100000, 3, 24), [0, 2], T.RED)]:
    for dt in (torch.float32, torch.bfloat16, torch.float64, torch.int32):
        x = (torch.rand(shape, device='cuda', generator=g) * 64).to(dt)
        assert T.plan_kind(x, axes)[0] == kind, (shape, axes, T.plan_kind(x, axes))
        want = x.double().sum(dim=axes)
        got = pb.sum(x, axis=axes, dtype='float32' if dt == torch.bfloat16 else None)
        tol = 0 if dt == torch.int32 else (1e-12 if dt == torch.float64 else 1e-5)
        assert bool(((got.double() - want).abs() <= tol * want.abs()).all()), (shape, axes, dt)
        assert torch.equal(pb.max(x, axis=axes), x.amax(dim=axes))
print('tiled ok')
"""
    env = dict(os.environ, B200PHI_NO_TWO_STAGE="1")
    r = subprocess.run([sys.executable, "-c", script], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "tiled ok" in r.stdout, r.stdout + r.stderr
