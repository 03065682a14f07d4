"""b200phi_reduce_noalloc, the entry point for hosts that cannot allocate device memory
where they call the library (no allocator linked, a frozen memory plan, an allocator lock
they must not take): its scratch is a fixed 4 MiB array inside the library.

The file name sorts last on purpose: if this test fails under `pytest -x`, every other
test has already run.  It must pass both alone in a fresh process and after the others.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_util as G
from paddle2_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _desc(x, out, op=L.SUM):
    d = L.ReduceDesc()
    d.op = op
    d.in_dtype = {torch.float32: L.FLOAT32, torch.bfloat16: L.BFLOAT16, torch.int8: L.INT8}[x.dtype]
    d.out_dtype = {torch.float32: L.FLOAT32, torch.bfloat16: L.BFLOAT16, torch.int64: L.INT64}[out.dtype]
    d.rank = x.dim()
    for i, s in enumerate(x.shape):
        d.dims[i] = s
    d.naxes = 1
    d.axes[0] = 0
    d.x = x.data_ptr()
    d.out = out.data_ptr()
    return d


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _with_caller_workspace(x, out, op=L.SUM):
    """The existing entry point, scratch supplied by the caller."""
    lib = L.lib()
    d = _desc(x, out, op)
    ws_bytes = lib.b200phi_reduce_workspace(C.byref(d))
    assert ws_bytes > 0, "the shape under test must be one that needs scratch"
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws_bytes
    L.check(lib.b200phi_reduce(C.byref(d), _stream()))
    torch.cuda.synchronize()
    return ws_bytes


@pytest.mark.parametrize("shape,dtype", [((40000, 1024), torch.float32), ((7, 3000, 512), torch.float32),
                                         ((30011, 768), torch.bfloat16)])
def test_matches_the_workspace_entry_point_twice_in_a_row(shape, dtype):
    lib = L.lib()
    g = torch.Generator(device="cuda").manual_seed(123)
    x = torch.rand(shape, device="cuda", generator=g).to(dtype)
    out_shape = shape[1:] if len(shape) == 2 else (shape[0], shape[2])
    axis = 0 if len(shape) == 2 else 1
    want = torch.empty(out_shape, device="cuda", dtype=dtype)
    d0 = _desc(x, want)
    d0.axes[0] = axis
    ws_bytes = lib.b200phi_reduce_workspace(C.byref(d0))
    assert 0 < ws_bytes <= L.NOALLOC_SCRATCH_BYTES
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    d0.workspace, d0.workspace_bytes = ws.data_ptr(), ws_bytes
    L.check(lib.b200phi_reduce(C.byref(d0), _stream()))
    tol = G.BF16_RTOL if dtype == torch.bfloat16 else G.FP32_RTOL
    for call in range(2):                      # the second call finds the first one's leftovers
        got = torch.full(out_shape, float("nan"), device="cuda", dtype=dtype)
        d1 = _desc(x, got)
        d1.axes[0] = axis                      # no workspace fields set: the library owns the scratch
        L.check(lib.b200phi_reduce_noalloc(C.byref(d1), _stream()))
        torch.cuda.synchronize()
        np.testing.assert_allclose(G.to_np(got), G.to_np(want), rtol=tol, err_msg=f"call {call}")
    # a different reduce in between must not disturb the next one (shared array, same stream)
    y = torch.rand(50000, 256, device="cuda", generator=g)
    oy = torch.empty(256, device="cuda")
    L.check(lib.b200phi_reduce_noalloc(C.byref(_desc(y, oy)), _stream()))
    got = torch.empty(out_shape, device="cuda", dtype=dtype)
    d1 = _desc(x, got)
    d1.axes[0] = axis
    L.check(lib.b200phi_reduce_noalloc(C.byref(d1), _stream()))
    torch.cuda.synchronize()
    np.testing.assert_allclose(G.to_np(got), G.to_np(want), rtol=tol)
    np.testing.assert_allclose(G.to_np(oy), G.to_np(y.double().sum(0).float()), rtol=G.FP32_RTOL)


def test_max_uses_the_same_path():
    lib = L.lib()
    x = torch.rand(60000, 640, device="cuda")
    want = torch.empty(640, device="cuda")
    _with_caller_workspace(x, want, L.MAX)
    got = torch.empty(640, device="cuda")
    L.check(lib.b200phi_reduce_noalloc(C.byref(_desc(x, got, L.MAX)), _stream()))
    torch.cuda.synchronize()
    assert torch.equal(got, want)


def test_long_rows_and_two_stage_shapes_use_the_built_in_scratch_too():
    """A full reduce of a long 1-D tensor (row cut across CTAs) and the per-channel sum of
    [N,C,HW] (row stage -> intermediate -> column stage) also need scratch; twice in a row each."""
    lib = L.lib()
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.rand(1 << 26, device="cuda", generator=g)
    want = torch.empty((), device="cuda")
    assert _with_caller_workspace(x, want) <= L.NOALLOC_SCRATCH_BYTES
    for call in range(2):
        got = torch.full((), float("nan"), device="cuda")
        L.check(lib.b200phi_reduce_noalloc(C.byref(_desc(x, got)), _stream()))
        torch.cuda.synchronize()
        assert torch.equal(got, want), f"call {call}"
    x = torch.rand(64, 96, 3136, device="cuda", generator=g)

    def two_axes(out):
        d = _desc(x, out)
        d.naxes = 2
        d.axes[0], d.axes[1] = 0, 2
        return d

    want = torch.empty(96, device="cuda")
    d0 = two_axes(want)
    ws_bytes = lib.b200phi_reduce_workspace(C.byref(d0))
    assert 0 < ws_bytes <= L.NOALLOC_SCRATCH_BYTES
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    d0.workspace, d0.workspace_bytes = ws.data_ptr(), ws_bytes
    L.check(lib.b200phi_reduce(C.byref(d0), _stream()))
    for call in range(2):
        got = torch.full((96,), float("nan"), device="cuda")
        L.check(lib.b200phi_reduce_noalloc(C.byref(two_axes(got)), _stream()))
        torch.cuda.synchronize()
        assert torch.equal(got, want), f"call {call}"


def test_refuses_what_does_not_fit_and_serves_what_needs_no_scratch():
    lib = L.lib()
    # int8 -> int64 column sum: 16 int64 partials per 16-byte vector, far more than 4 MiB
    x = torch.zeros(4096, 65536, device="cuda", dtype=torch.int8)
    out = torch.empty(65536, device="cuda", dtype=torch.int64)
    d = _desc(x, out)
    need = lib.b200phi_reduce_workspace(C.byref(d))
    assert need > L.NOALLOC_SCRATCH_BYTES
    rc = lib.b200phi_reduce_noalloc(C.byref(d), _stream())
    assert rc == 4                                           # B200PHI_WORKSPACE_TOO_SMALL
    msg = lib.b200phi_last_error().decode()
    assert str(need) in msg and str(L.NOALLOC_SCRATCH_BYTES) in msg
    # a last-axis reduce needs no scratch: served all the same, the array stays untouched
    x = torch.rand(64, 4096, device="cuda")
    out = torch.empty(64, device="cuda")
    d = _desc(x, out)
    d.axes[0] = 1
    L.check(lib.b200phi_reduce_noalloc(C.byref(d), _stream()))
    torch.cuda.synchronize()
    np.testing.assert_allclose(G.to_np(out), G.to_np(x.double().sum(1).float()), rtol=G.FP32_RTOL)
