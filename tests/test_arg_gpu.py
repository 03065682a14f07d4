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
