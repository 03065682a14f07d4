"""The C-ABI library loads on a CPU-only box, exports every symbol include/b200phi.h
declares, and its host-side planning (no kernel launch) reports errors the way the
reference does.  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200phi.h")).read()
    return sorted(set(re.findall(r"B200PHI_API\s+[\w\s\*]+?\b(b200phi_\w+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    from paddle2_b200 import _lib
    assert _declared_symbols() == sorted(_lib.EXPORTS)


def test_library_exports_every_declared_symbol():
    from paddle2_b200 import _lib
    lib = _lib.lib()
    for name in _declared_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.b200phi_version()


def test_struct_layout_matches_header():
    """ctypes mirrors of the POD descriptors must have the C sizes (checked by compiling
    a tiny C program against the header)."""
    import subprocess
    import tempfile
    from paddle2_b200 import _lib
    src = r'''
#include <stdio.h>
#include "b200phi.h"
int main(void){printf("%zu %zu %zu %zu %zu\n", sizeof(b200phi_reduce_desc), sizeof(b200phi_argreduce_desc),
 sizeof(b200phi_rowstats_desc), sizeof(b200phi_tensor_view), sizeof(b200phi_elementwise_desc)); return 0;}
'''
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(td, "s")
        subprocess.run(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe], check=True)
        sizes = [int(v) for v in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(_lib.ReduceDesc), C.sizeof(_lib.ArgReduceDesc), C.sizeof(_lib.RowStatsDesc),
                     C.sizeof(_lib.TensorView), C.sizeof(_lib.ElementwiseDesc)]


def _reduce_desc(dims, axes, op=0, in_dtype=10, out_dtype=0, reduce_all=0):
    from paddle2_b200 import _lib
    d = _lib.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = op, in_dtype, out_dtype, len(dims)
    for i, s in enumerate(dims):
        d.dims[i] = s
    d.naxes = len(axes)
    for i, a in enumerate(axes):
        d.axes[i] = a
    d.reduce_all = reduce_all
    return d


def test_planner_errors_match_reference_messages():
    from paddle2_b200 import _lib
    lib = _lib.lib()
    # empty tensor: reduce_function.h:1052-1055
    d = _reduce_desc([0, 3], [0])
    assert lib.b200phi_reduce(C.byref(d), None) == 1
    assert b"Tensor need be reduced must not empty." in lib.b200phi_last_error()
    # axis out of range: reduce_function.h:112-118
    d = _reduce_desc([2, 3], [2])
    assert lib.b200phi_reduce(C.byref(d), None) == 1
    assert b"should be in the range" in lib.b200phi_last_error()
    # unknown dtype / unsupported pairs
    d = _reduce_desc([2, 3], [1], in_dtype=14)                       # PSTRING
    assert lib.b200phi_reduce(C.byref(d), None) == 2
    d = _reduce_desc([2, 3], [1], op=0, in_dtype=17)                 # float8: max / amax only
    assert lib.b200phi_reduce(C.byref(d), None) == 2
    assert b"float8" in lib.b200phi_last_error()
    d = _reduce_desc([2, 3], [1], op=2, in_dtype=12)                 # complex: sum / mean only
    assert lib.b200phi_reduce(C.byref(d), None) == 2
    d = _reduce_desc([2, 3], [1], op=5, in_dtype=10, out_dtype=10)   # any writes BOOL
    assert lib.b200phi_reduce(C.byref(d), None) == 1
    # out_dtype != x.dtype where the cast cannot be folded into the load: the reference's Cast pre-pass,
    # scratch = the tensor in the output dtype (+ what the reduce itself needs)
    d = _reduce_desc([4096, 512], [0], op=0, in_dtype=10, out_dtype=16)
    assert lib.b200phi_reduce_plan_kind(C.byref(d)) == 12
    assert lib.b200phi_reduce_workspace(C.byref(d)) >= 4096 * 512 * 2
    d = _reduce_desc([4096, 512], [0], op=2, in_dtype=10, out_dtype=16)   # max fp32 -> bf16: converts at the store
    assert lib.b200phi_reduce_plan_kind(C.byref(d)) == 3
    # complex: planned as float lanes (dims + [2])
    d = _reduce_desc([4096, 512], [0], op=0, in_dtype=12)
    assert lib.b200phi_reduce_plan_kind(C.byref(d)) == 3
    with pytest.raises(_lib.B200PhiError):
        _lib.check(1)


def test_short_rows_take_the_staged_plan_by_the_measured_limits():
    """kRowStaged (14): rows of 128 bulk-copied into shared memory, a thread per row
    (rows_staged.cuh).  Selection is host logic (reduce_plan.h: staged_rows_ok)."""
    from paddle2_b200 import _lib
    lib = _lib.lib()
    STAGED = 14

    def kind(rows, cols, dt, x=0, op=0):
        d = _reduce_desc([rows, cols], [1], op=op, in_dtype=dt)
        d.x = x
        return lib.b200phi_reduce_plan_kind(C.byref(d)), lib.b200phi_reduce_workspace(C.byref(d))

    A = 1 << 20                                            # a 16-byte aligned address, B = A + 2 is not
    assert kind(100000, 101, 16, A) == (STAGED, 0)          # bf16 202-byte rows, ragged
    assert kind(100000, 101, 16, A + 2) == (STAGED, 0)      # ... whatever the base alignment
    assert kind(100000, 21, 10, A)[0] == STAGED             # fp32 84-byte rows
    assert kind(100000, 10, 10, A)[0] == STAGED             # fp32 40 bytes
    assert kind(100000, 7, 10, A)[0] != STAGED              # plain reductions under 32 bytes (fp32): sub-warp kernel
    assert kind(100000, 7, 9, A)[0] != STAGED               # ... under 64 bytes for 8-byte elements
    assert kind(100000, 301, 16, A)[0] != STAGED            # 602 bytes: beyond the 16-bit limit (512)
    assert kind(100000, 201, 10, A)[0] == STAGED            # fp32 804 bytes: within 832
    assert kind(100000, 251, 10, A)[0] != STAGED
    assert kind(1000, 101, 16, A)[0] != STAGED              # too few rows to fill the machine
    # rows that are whole 16-byte vectors: only where the vector kernels are weak, and only from a
    # 16-byte aligned base (they are read with 16-byte shared loads)
    assert kind(100000, 32, 10, A)[0] != STAGED             # fp32 128 bytes = 8 vectors: row_tiny_kernel
    assert kind(100000, 40, 10, A)[0] == STAGED             # 10 vectors: not a power of two
    assert kind(100000, 40, 10, A + 4)[0] != STAGED
    assert kind(100000, 64, 10, A)[0] == STAGED             # 256 bytes: from 192 bytes on for 4-byte elements
    assert kind(100000, 96, 16, A)[0] == STAGED             # bf16 192 bytes = 12 vectors
    assert kind(100000, 128, 16, A)[0] != STAGED            # bf16 256 bytes = 16 vectors
    # every plan of this family needs no scratch, whatever the pointer
    assert kind(100000, 40, 10, 0)[1] == 0


def test_workspace_query_is_pure_host_logic():
    from paddle2_b200 import _lib
    lib = _lib.lib()
    # C1: row reduce, no scratch
    assert lib.b200phi_reduce_workspace(C.byref(_reduce_desc([32, 1024, 4096], [2]))) == 0
    # C5: axis-0 reduce of [524288, 4096] splits rows across CTAs -> partials + tickets
    ws = lib.b200phi_reduce_workspace(C.byref(_reduce_desc([524288, 4096], [0])))
    assert ws > 4096 * 4 and ws % 4 == 0
    # full reduction of 2^31 elements (the numel > INT32_MAX regime, kps/reduce_kernel.cu:196)
    ws = lib.b200phi_reduce_workspace(C.byref(_reduce_desc([2 ** 31], [0])))
    assert 0 < ws < (1 << 20)


def test_arg_workspace_query_is_pure_host_logic():
    from paddle2_b200 import _lib
    lib = _lib.lib()

    def desc(dims, axis=None, dt=10):
        d = _lib.ArgReduceDesc()
        d.op, d.in_dtype, d.index_dtype, d.rank = 0, dt, 9, len(dims)
        for i, s in enumerate(dims):
            d.dims[i] = s
        d.flatten = 1 if axis is None else 0
        d.axis = 0 if axis is None else axis
        return d

    # C3: many rows over the last axis -> a group of threads per row, no scratch
    assert lib.b200phi_argreduce_workspace(C.byref(desc([8192, 32000], 1))) == 0
    # flatten (paddle.argmax's default axis=None) of 2^30 bf16: split across CTAs
    ws = lib.b200phi_argreduce_workspace(C.byref(desc([1 << 30], None, 16)))
    assert 0 < ws < (1 << 20)
    # the sharded argmax's local step: axis 0 of [524288, 4096] fp32 -> (value, index) partials per split
    ws = lib.b200phi_argreduce_workspace(C.byref(desc([524288, 4096], 0)))
    assert ws >= 2 * 4096 * 12 and ws < (64 << 20)
    # bad descriptors answer 0 and the call reports the reference's wording
    d = desc([0, 4], 1)
    assert lib.b200phi_argreduce_workspace(C.byref(d)) == 0
    assert lib.b200phi_argreduce(C.byref(d), None) == 1
    assert b"numel must > 0" in lib.b200phi_last_error()
    d = desc([3, 4], 2)
    assert lib.b200phi_argreduce(C.byref(d), None) == 1
    assert b"axis out of range" in lib.b200phi_last_error()


def test_sharded_workspace_adds_the_wide_partial_only_for_16bit_sums():
    from paddle2_b200 import _lib
    lib = _lib.lib()
    base = lib.b200phi_reduce_workspace(C.byref(_reduce_desc([524288, 4096], [0])))
    assert lib.b200phi_sharded_reduce_workspace(C.byref(_reduce_desc([524288, 4096], [0]))) == base
    b16 = lib.b200phi_reduce_workspace(C.byref(_reduce_desc([524288, 4096], [0], in_dtype=16)))
    w16 = lib.b200phi_sharded_reduce_workspace(C.byref(_reduce_desc([524288, 4096], [0], in_dtype=16)))
    assert w16 >= b16 + 4096 * 4
    # max keeps the dtype: no rounding, no wide partial
    m16 = lib.b200phi_sharded_reduce_workspace(C.byref(_reduce_desc([524288, 4096], [0], op=2, in_dtype=16)))
    assert m16 == lib.b200phi_reduce_workspace(C.byref(_reduce_desc([524288, 4096], [0], op=2, in_dtype=16)))


def test_workspace_query_ignores_the_data_pointer():
    from paddle2_b200 import _lib
    lib = _lib.lib()
    for dims, axes, dt in (([40000, 1024], [0], 10), ([30011, 768], [0], 16), ([7, 3000, 512], [1], 10)):
        d = _reduce_desc(dims, axes, in_dtype=dt)
        a = lib.b200phi_reduce_workspace(C.byref(d))
        d.x = 0x7f0000000010
        b = lib.b200phi_reduce_workspace(C.byref(d))
        d.x = 0x7f0000000012
        c = lib.b200phi_reduce_workspace(C.byref(d))
        assert a == b == c > 0


def test_elementwise_plan_errors():
    from paddle2_b200 import _lib
    lib = _lib.lib()
    d = _lib.ElementwiseDesc()
    d.binary, d.act, d.dtype, d.arity, d.axis = 0, 0, 10, 2, -1
    d.in_[0].rank = 2
    d.in_[0].dims[0], d.in_[0].dims[1] = 2, 3
    d.in_[1].rank = 2
    d.in_[1].dims[0], d.in_[1].dims[1] = 2, 4
    d.out_rank = 2
    d.out_dims[0], d.out_dims[1] = 2, 4
    assert lib.b200phi_elementwise(C.byref(d), None) == 1
    assert b"Broadcast dimension mismatch" in lib.b200phi_last_error()


def test_no_cpu_fallback():
    """The product path refuses host tensors instead of silently computing on the CPU."""
    import torch
    import paddle2_b200 as pb
    with pytest.raises(RuntimeError, match="no CPU path"):
        pb.sum(torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="no CPU path"):
        pb.add(torch.ones(4), torch.ones(4))


def test_product_code_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under paddle2_b200/ may reference it."""
    bad = []
    for base, _dirs, files in os.walk(os.path.join(ROOT, "paddle2_b200")):
        if "build" in base.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")) or f == "Makefile":
                text = open(os.path.join(base, f), errors="ignore").read()
                if "pyoracle" in text or "phi_oracle" in text or "libphi_ref" in text:
                    bad.append(os.path.join(base, f))
    assert not bad, bad
