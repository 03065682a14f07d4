"""Out-of-bounds writes: every output (and scratch) is carved out of a larger buffer whose
margins hold a sentinel; after the call the margins must be untouched.  (compute-sanitizer is
closed on this GPU pool, so the suite carries its own guard.)"""
import ctypes as C

import numpy as np
import pytest
import torch

import paddle2_b200 as pb
from paddle2_b200 import _lib as L

pytestmark = pytest.mark.gpu
GUARD = 4096  # bytes on each side


def _guarded(shape, dtype):
    n = int(np.prod(shape)) if len(shape) else 1
    esize = torch.empty((), dtype=dtype).element_size()
    raw = torch.full((2 * GUARD + n * esize,), 0xA5, dtype=torch.uint8, device="cuda")
    view = raw[GUARD:GUARD + n * esize].view(dtype).view(shape)
    return raw, view


def _intact(raw, nbytes):
    return bool((raw[:GUARD] == 0xA5).all()) and bool((raw[GUARD + nbytes:] == 0xA5).all())


@pytest.mark.parametrize("shape,axis", [((1027, 513), [1]), ((1027, 513), [0]), ((40001, 96), [0]), ((3, 99991), [1]),
                                        ((6, 5, 40, 7), [1, 3]), ((100003, 3), [1]), ((999983,), None)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64])
def test_reduce_outputs_and_scratch_stay_in_bounds(shape, axis, dtype):
    x = (torch.rand(shape, device="cuda") * 100).to(dtype)
    want = pb.sum(x, axis=axis)
    raw, out = _guarded(tuple(want.shape), want.dtype)
    d = L.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = L.SUM, pb.ops._phi_dtype(x.dtype), pb.ops._phi_dtype(want.dtype), x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    axes = [] if axis is None else axis
    d.naxes = len(axes)
    for i, a in enumerate(axes):
        d.axes[i] = a
    d.x, d.out = x.data_ptr(), out.data_ptr()
    lib = L.lib()
    ws_bytes = lib.b200phi_reduce_workspace(C.byref(d))
    wraw = None
    if ws_bytes:
        pad = (256 - GUARD % 256) % 256
        wraw = torch.full((2 * GUARD + pad + ws_bytes + 256,), 0xA5, dtype=torch.uint8, device="cuda")
        base = wraw.data_ptr() + GUARD
        off = (256 - base % 256) % 256
        d.workspace, d.workspace_bytes = base + off, ws_bytes
    L.check(lib.b200phi_reduce(C.byref(d), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    assert _intact(raw, want.numel() * want.element_size())
    if wraw is not None:
        assert bool((wraw[:GUARD] == 0xA5).all())
        end = GUARD + off + ws_bytes
        assert bool((wraw[end:] == 0xA5).all())


@pytest.mark.parametrize("xs,ys", [((257, 1031), (1031,)), ((64, 3, 501), (3, 1)), ((1000, 7), (1000, 1)), ((10, 3, 1, 4), (10, 1, 12, 1))])
def test_elementwise_output_stays_in_bounds(xs, ys):
    x = torch.rand(xs, device="cuda")
    y = torch.rand(ys, device="cuda")
    want = pb.fused_elemwise_add_activation(x, y, ["relu", "elementwise_add"])
    raw, out = _guarded(tuple(want.shape), torch.float32)
    e = L.ElementwiseDesc()
    e.binary, e.act, e.compound, e.dtype, e.arity, e.axis = L.ADD, L.ACT_RELU, 0, L.FLOAT32, 2, -1
    for i, t in enumerate((x, y)):
        e.in_[i].rank = t.dim()
        for j, s in enumerate(t.shape):
            e.in_[i].dims[j] = s
        e.in_[i].data = t.data_ptr()
    e.out_rank = want.dim()
    for j, s in enumerate(want.shape):
        e.out_dims[j] = s
    e.out = out.data_ptr()
    L.check(L.lib().b200phi_elementwise(C.byref(e), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    assert _intact(raw, want.numel() * 4)


def test_arg_and_rowstats_outputs_stay_in_bounds():
    x = torch.randn(1003, 4099, device="cuda")
    want = pb.argmax(x, axis=1)
    raw, out = _guarded((1003,), torch.int64)
    a = L.ArgReduceDesc()
    a.op, a.in_dtype, a.index_dtype, a.rank = L.ARGMAX, L.FLOAT32, L.INT64, 2
    a.dims[0], a.dims[1] = 1003, 4099
    a.axis = 1
    a.x, a.out = x.data_ptr(), out.data_ptr()
    L.check(L.lib().b200phi_argreduce(C.byref(a), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert torch.equal(out, want) and _intact(raw, 1003 * 8)
    xb = torch.randn(999, 8192, device="cuda").bfloat16()
    m, v = pb.row_mean_var(xb)
    rawm, om = _guarded((999,), torch.bfloat16)
    rawv, ov = _guarded((999,), torch.bfloat16)
    r = L.RowStatsDesc()
    r.in_dtype, r.out_dtype, r.rows, r.cols, r.unbiased = L.BFLOAT16, L.BFLOAT16, 999, 8192, 1
    r.x, r.mean, r.var = xb.data_ptr(), om.data_ptr(), ov.data_ptr()
    L.check(L.lib().b200phi_rowstats(C.byref(r), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert torch.equal(om, m) and torch.equal(ov, v)
    assert _intact(rawm, 999 * 2) and _intact(rawv, 999 * 2)


def _fill_views(desc_views, tensors):
    for view, t in zip(desc_views, tensors):
        view.rank = t.dim()
        for j, s in enumerate(t.shape):
            view.dims[j] = s
        view.data = t.data_ptr()


@pytest.mark.parametrize("xs,ys", [((257, 1031), (1031,)), ((64, 3, 501), (3, 1)), ((1000, 7), (1000, 1)),
                                   ((10, 3, 1, 4), (10, 1, 12, 1)), ((33, 4099), (33, 4099))])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64, torch.uint8])
def test_map_outputs_stay_in_bounds(xs, ys, dtype):
    """compare writes a packed group of bool bytes per input vector (2..16 bytes): odd widths and
    unaligned rows must not spill; same for the ternary functor's same-dtype output"""
    x = (torch.rand(xs, device="cuda") * 9).to(dtype)
    y = (torch.rand(ys, device="cuda") * 9).to(dtype)
    want = pb.less_equal(x, y)
    raw, out = _guarded(tuple(want.shape), torch.bool)
    d = L.MapDesc()
    d.op, d.in_dtype, d.out_dtype, d.arity, d.axis = L.MAP_OPS["less_equal"], pb.ops._phi_dtype(dtype), 0, 2, -1
    _fill_views(d.in_, (x, y))
    d.out_rank = want.dim()
    for j, s in enumerate(want.shape):
        d.out_dims[j] = s
    d.out = out.data_ptr()
    L.check(L.lib().b200phi_map(C.byref(d), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    assert _intact(raw, want.numel())
    if dtype in (torch.float32, torch.bfloat16, torch.int64):
        g = (torch.rand(tuple(want.shape), device="cuda") * 3).to(dtype)
        want3 = pb.ops._map("max_grad_x", [x.expand(want.shape).contiguous(), y.expand(want.shape).contiguous(), g])
        raw3, out3 = _guarded(tuple(want.shape), dtype)
        d3 = L.MapDesc()
        d3.op, d3.in_dtype, d3.out_dtype, d3.arity, d3.axis = L.MAP_OPS["max_grad_x"], pb.ops._phi_dtype(dtype), 0, 3, -1
        _fill_views(d3.in_, (x, y, g))
        d3.out_rank = want.dim()
        for j, s in enumerate(want.shape):
            d3.out_dims[j] = s
        d3.out = out3.data_ptr()
        L.check(L.lib().b200phi_map(C.byref(d3), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        torch.cuda.synchronize()
        assert torch.equal(out3, want3)
        assert _intact(raw3, want3.numel() * want3.element_size())


@pytest.mark.parametrize("rows,cols", [(1027, 513), (3, 99991), (100003, 3), (64, 32000), (5, 4099)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_logsumexp_and_mixed_add_stay_in_bounds(rows, cols, dtype):
    x = torch.randn((rows, cols), device="cuda").to(dtype)
    want = pb.ops.logsumexp(x, -1)
    raw, out = _guarded((rows,), dtype)
    d = L.LogsumexpDesc()
    d.dtype, d.rows, d.cols, d.x, d.out = pb.ops._phi_dtype(dtype), rows, cols, x.data_ptr(), out.data_ptr()
    L.check(L.lib().b200phi_logsumexp(C.byref(d), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    assert _intact(raw, rows * want.element_size())
    if dtype == torch.bfloat16:
        acc = torch.randn((rows, cols), device="cuda")
        want2 = acc + x.float()
        raw2, out2 = _guarded((rows, cols), torch.float32)
        m = L.MixedAddDesc()
        m.y_dtype = L.BFLOAT16
        _fill_views((m.x, m.y), (acc, x))
        m.out_rank = 2
        m.out_dims[0], m.out_dims[1] = rows, cols
        m.out = out2.data_ptr()
        L.check(L.lib().b200phi_mixed_add(C.byref(m), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        torch.cuda.synchronize()
        assert torch.equal(out2, want2)
        assert _intact(raw2, rows * cols * 4)
