"""argmax / argmin parity: indices are bit-exact, ties go to the lowest index."""
import numpy as np
import pytest
import torch

import cases
import gpu_util as G
import paddle2_b200 as pb

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("i", range(len(cases.ARG_CASES)))
def test_replay_reference_cases(oracle, i):
    op, dt, shape, axis, keepdim, idt, src = cases.ARG_CASES[i]
    x = (1000 * np.random.default_rng(2000 + i).random(shape)).astype(dt)
    want = oracle.argreduce(op, x, axis, keepdim, np.dtype(idt))
    got = G.to_np(getattr(pb, op)(G.to_dev(x), axis=axis, keepdim=keepdim, dtype=idt))
    assert got.dtype == want.dtype and got.shape == want.shape, src
    assert np.array_equal(got, want), src


@pytest.mark.parametrize("shape,axis", [((3000, 10), 1), ((2400, 1027), 1), ((64, 70001), 1),
                                        ((1 << 20,), None), ((50, 300, 33), 1), ((4000, 9), 0),
                                        ((7, 9, 11, 13), 2), ((5, 64, 256), 1)])
@pytest.mark.parametrize("dt", ["float32", "float64", "int32", "int64", "uint8", "int16"])
def test_ties_everywhere(oracle, shape, axis, dt):
    """Few distinct values => many ties at every combine level (thread, warp, CTA)."""
    rng = np.random.default_rng(abs(hash((shape, dt))) % (2 ** 31))
    x = rng.integers(0, 5, size=shape).astype(dt)
    for op in ("argmax", "argmin"):
        want = oracle.argreduce(op, x, axis)
        got = G.to_np(getattr(pb, op)(G.to_dev(x), axis=axis))
        assert np.array_equal(got, want), (op, shape, dt)


def test_bf16_and_fp16_indices(oracle):
    rng = np.random.default_rng(9)
    x32 = G.bf16_round(rng.standard_normal((512, 32000)).astype(np.float32))  # natural ties
    xt = torch.from_numpy(x32).cuda().bfloat16()
    for op in ("argmax", "argmin"):
        assert np.array_equal(G.to_np(getattr(pb, op)(xt, axis=-1)), oracle.argreduce(op, x32, -1))
    idx, val = pb.argmax_with_value(xt, axis=-1)
    assert np.array_equal(G.to_np(idx), oracle.argreduce("argmax", x32, -1))
    assert np.array_equal(G.to_np(val), oracle.reduce("max", x32, [-1]))
    xh = torch.from_numpy(x32).cuda().half()
    assert np.array_equal(G.to_np(pb.argmax(xh, axis=0)),
                          oracle.argreduce("argmax", xh.float().cpu().numpy(), 0))


def test_identity_rows_and_extremes(oracle):
    i64 = np.iinfo(np.int64)
    x = np.array([[1, 3, 3, 2], [i64.min] * 4, [5, 1, 5, 5], [i64.max, 0, i64.max, 1]], dtype=np.int64)
    assert np.array_equal(G.to_np(pb.argmax(G.to_dev(x), axis=1)), [1, 0, 0, 0])
    assert np.array_equal(G.to_np(pb.argmin(G.to_dev(x), axis=1)), oracle.argreduce("argmin", x, 1))
    f = np.full((3, 1000), -np.inf, dtype=np.float32)
    assert np.array_equal(G.to_np(pb.argmax(G.to_dev(f), axis=1)), oracle.argreduce("argmax", f, 1))


def test_int32_indices_keepdim_flatten(oracle):
    x = np.random.default_rng(1).random((6, 7, 8)).astype(np.float32)
    got = pb.argmax(G.to_dev(x), axis=1, keepdim=True, dtype="int32")
    assert got.dtype == torch.int32 and tuple(got.shape) == (6, 1, 8)
    assert np.array_equal(G.to_np(got), oracle.argreduce("argmax", x, 1, True, np.int32))
    got = pb.argmin(G.to_dev(x), keepdim=True)
    assert tuple(got.shape) == (1, 1, 1) and int(got) == int(np.argmin(x))
    with pytest.raises(ValueError):
        pb.argmax(torch.empty(0, device="cuda"))


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16, torch.int16])
@pytest.mark.parametrize("rows,cols", [(64, 257), (333, 1001), (129, 1031), (65, 2049), (70, 3001), (67, 4097), (66, 4103)])
def test_16bit_unaligned_rows_warp_peel(rows, cols, dt):
    """16-bit rows that are not whole aligned vectors (arg_warp_peel_kernel): many ties (small value
    range) so the first occurrence matters, head and tail elements included; also from a base one
    element past a 16-byte boundary"""
    import numpy as np
    import paddle2_b200 as pb
    gen = torch.Generator().manual_seed(rows * 5 + cols)
    flat = torch.randint(-6, 7, (rows * cols + 1,), generator=gen).to(dt)
    fd = flat.cuda()
    for off in (0, 1):
        x = flat[off:off + rows * cols].view(rows, cols)
        xd = fd[off:off + rows * cols].view(rows, cols)
        ref = x.float().numpy()
        assert np.array_equal(pb.argmax(xd, axis=-1).cpu().numpy(), np.argmax(ref, axis=-1))
        assert np.array_equal(pb.argmin(xd, axis=-1).cpu().numpy(), np.argmin(ref, axis=-1))
        idx, val = pb.argmax_with_value(xd, axis=-1)
        assert np.array_equal(idx.cpu().numpy(), np.argmax(ref, axis=-1))
        assert torch.equal(val.cpu(), x.amax(-1))
    # the extreme sits in the head / in the tail of the row
    x = torch.zeros(rows, cols, dtype=dt)
    x[:, 0] = 3
    x[1::2, 0] = 0
    x[1::2, cols - 1] = 5
    xd = x.cuda()
    assert np.array_equal(pb.argmax(xd, axis=-1).cpu().numpy(), np.argmax(x.float().numpy(), axis=-1))


# ---- split plans: flatten / few long rows (arg_row_split_kernel), non-last axis (arg_col_split_kernel)
def _arg_plan(x, axis):
    """(workspace bytes the C ABI asks for) -- non-zero means the reduced extent is split across CTAs."""
    import ctypes as C
    from paddle2_b200 import _lib as L
    d = L.ArgReduceDesc()
    d.op, d.in_dtype, d.index_dtype, d.rank = L.ARGMAX, pb.ops._phi_dtype(x.dtype), L.INT64, x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.flatten = 1 if axis is None else 0
    d.axis = 0 if axis is None else axis
    return L.lib().b200phi_argreduce_workspace(C.byref(d)), d


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16, torch.int64, torch.uint8, torch.float64])
@pytest.mark.parametrize("n,off", [(1 << 22, 0), ((1 << 22) + 12345, 0), ((1 << 21) + 3, 1)])
def test_flatten_is_split_across_ctas_ties_at_every_boundary(dt, n, off):
    """paddle.argmax's default axis=None: one long row.  Few distinct values, and the extreme planted on
    BOTH sides of every split boundary (so equal maxima sit in different CTAs): the lowest index wins."""
    gen = torch.Generator(device="cuda").manual_seed(n % 1000 + off)
    flat = torch.randint(0, 4, (n + off,), device="cuda", generator=gen).to(dt)
    x = flat[off:]
    ws, _ = _arg_plan(x, None)
    assert ws > 0, "this size must take the split plan"
    ref = x.float().cpu().numpy() if dt != torch.int64 else x.cpu().numpy()
    assert int(pb.argmax(x)) == int(np.argmax(ref)) and int(pb.argmin(x)) == int(np.argmin(ref))
    # plant the maximum 9 at both sides of every possible segment boundary of a few split counts
    y = x.clone()
    first = None
    for splits in (37, 148, 592):
        seg = -(-n // splits)
        for b in range(seg, n, seg):
            for j in (b - 1, b):
                y[j] = 9
                first = j if first is None else min(first, j)
    idx, val = pb.argmax_with_value(y)
    assert int(idx) == first and float(val) == 9.0
    y2 = x.clone()
    y2[n - 1] = 9            # the very last element, alone
    assert int(pb.argmax(y2)) == n - 1
    y2[0] = 9
    assert int(pb.argmax(y2)) == 0
    got32 = pb.argmax(y, dtype="int32")
    assert got32.dtype == torch.int32 and int(got32) == first


def test_flatten_all_identity_answers_zero():
    for dt, lo, hi in ((torch.float32, -3.4028234663852886e38, 3.4028234663852886e38),
                       (torch.int64, -(2 ** 63), 2 ** 63 - 1)):
        x = torch.full((1 << 22,), lo, device="cuda", dtype=dt)
        assert _arg_plan(x, None)[0] > 0
        assert int(pb.argmax(x)) == 0
        x = torch.full((1 << 22,), hi, device="cuda", dtype=dt)
        assert int(pb.argmin(x)) == 0
    f = torch.full((3, 1 << 21), float("-inf"), device="cuda")
    assert pb.argmax(f, axis=1).tolist() == [0, 0, 0]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16, torch.float16, torch.int64, torch.int32,
                                torch.int16, torch.uint8, torch.float64])
@pytest.mark.parametrize("shape,axis", [((50000, 512), 0), ((30011, 257), 0), ((7, 9000, 96), 1),
                                        ((3, 4001, 5), 1), ((2, 70000, 1, 2), 1), ((100003, 8), 0)])
def test_non_last_axis_split_with_ties(oracle, dt, shape, axis):
    """[pre, n, post] with n split over grid.y: vector path (post a whole number of 16-byte vectors),
    scalar path (post 257, 5, 2), ties across split boundaries, optional value output."""
    gen = torch.Generator(device="cuda").manual_seed(sum(shape) + axis)
    x = torch.randint(0, 5, shape, device="cuda", generator=gen).to(dt)
    ref = x.cpu().numpy() if dt in (torch.int64, torch.int32, torch.int16, torch.uint8, torch.float64) \
        else x.float().cpu().numpy()
    for op in ("argmax", "argmin"):
        want = oracle.argreduce(op, ref, axis)
        assert np.array_equal(G.to_np(getattr(pb, op)(x, axis=axis)), want), (op, shape, dt)
    idx, val = pb.argmin_with_value(x, axis=axis, keepdim=True)
    assert np.array_equal(G.to_np(idx), oracle.argreduce("argmin", ref, axis, True))
    assert np.array_equal(G.to_np(val).astype(np.float64), ref.min(axis=axis, keepdims=True).astype(np.float64))
    # one odd-offset base: the vector path must fall back cleanly
    if x.dim() == 2 and dt == torch.float32:
        flat = torch.cat([torch.zeros(1, device="cuda"), x.flatten()])
        xo = flat[1:].view(shape)
        assert np.array_equal(G.to_np(pb.argmax(xo, axis=axis)), oracle.argreduce("argmax", ref, axis))


def test_axis0_of_a_large_shard_matches_torch():
    """The local step of the sharded argmax: fp32 [65536*2, 4096] axis 0 (1 GiB)."""
    gen = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(131072, 4096, device="cuda", generator=gen)
    idx, val = pb.argmax_with_value(x, axis=0)
    tv, ti = torch.max(x, dim=0)
    assert torch.equal(val, tv)
    assert torch.equal(x.gather(0, idx[None, :])[0], tv)          # the index points at the maximum
    xb = x.bfloat16()                                              # natural ties: first occurrence
    idx = pb.argmax(xb, axis=0)
    m = xb.max(dim=0).values
    firsts = (xb == m[None, :]).to(torch.uint8).argmax(dim=0)
    assert torch.equal(idx, firsts)


def test_split_plans_refuse_a_missing_workspace():
    import ctypes as C
    from paddle2_b200 import _lib as L
    x = torch.rand(1 << 22, device="cuda")
    out = torch.empty((), device="cuda", dtype=torch.int64)
    ws, d = _arg_plan(x, None)
    d.x, d.out = x.data_ptr(), out.data_ptr()
    rc = L.lib().b200phi_argreduce(C.byref(d), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 4 and str(ws).encode() in L.lib().b200phi_last_error()
