"""Host-side rules that run before the kernel in Paddle (InferMeta) and that the
python mirror re-states: output shapes, dtype promotion, axis checks, broadcasting."""
import numpy as np
import pytest
import torch

import cases
from paddle2_b200 import dist as pdist
from paddle2_b200 import ops


def test_reduce_out_shape_rules():
    # ReduceInferDim, infermeta/unary.cc:3631-3700
    assert ops._norm_axes(None, 3) == ([0, 1, 2], True)
    assert ops._norm_axes([], 3) == ([0, 1, 2], True)
    assert ops._norm_axes([-1], 3) == ([2], False)
    assert ops._norm_axes([0, 1, 2], 3) == ([0, 1, 2], True)
    assert ops._norm_axes([-1, 2], 3) == ([2], False)
    assert ops._norm_axes(0, 0) == ([], True)
    with pytest.raises(ValueError):
        ops._norm_axes([3], 3)
    with pytest.raises(ValueError):
        ops._norm_axes([1], 0)
    assert ops._reduce_out_shape([5, 6, 10], [1], False, True) == [5, 1, 10]
    assert ops._reduce_out_shape([5, 6, 10], [1], False, False) == [5, 10]
    assert ops._reduce_out_shape([5, 6, 10], [0, 1, 2], True, True) == [1, 1, 1]
    assert ops._reduce_out_shape([5, 6, 10], [0, 1, 2], True, False) == []


def test_sum_dtype_promotion():
    # SumRawInferMeta, infermeta/unary.cc:4698-4707
    assert ops._sum_out_dtype(torch.int32, None) == torch.int64
    assert ops._sum_out_dtype(torch.bool, None) == torch.int64
    assert ops._sum_out_dtype(torch.float32, None) == torch.float32
    assert ops._sum_out_dtype(torch.bfloat16, None) == torch.bfloat16
    assert ops._sum_out_dtype(torch.int32, torch.float32) == torch.float32


@pytest.mark.parametrize("xs,ys,axis,src", cases.ADD_CASES + cases.FUSED_SHAPES)
def test_broadcast_shape_matches_numpy_layout(xs, ys, axis, src, oracle):
    got = ops.broadcast_shape(xs, ys, axis)
    want, xp, yp = oracle.broadcast_out_shape(xs, ys, axis)
    assert got == want, src
    assert list(np.broadcast_shapes(tuple(xp), tuple(yp))) == want


def test_broadcast_shape_errors():
    with pytest.raises(ValueError, match="Broadcast dimension mismatch"):
        ops.broadcast_shape([2, 3], [2, 4])
    with pytest.raises(ValueError):
        ops.broadcast_shape([2, 3], [2, 3], axis=1)  # same rank: axis must be -1 or 0


def test_argmax_argument_checks():
    # test_arg_min_max_v2_op.py:329-367 error cases (checked before any device work)
    x = torch.zeros(3, 4)
    with pytest.raises((ValueError, TypeError, RuntimeError)):
        ops.argmax(x, dtype="float32")
    with pytest.raises((ValueError, TypeError, RuntimeError)):
        ops.argmax(x, dtype=None)
    with pytest.raises((ValueError, TypeError, RuntimeError)):
        ops.argmax(x, axis=1.5)


def test_fused_functor_list_parsing():
    x = torch.zeros(2, 2)
    with pytest.raises(ValueError):
        ops.fused_elemwise_activation(x, x, ["relu"])
    with pytest.raises(ValueError):
        ops.fused_elemwise_activation(x, x, ["relu", "tanh"])
    with pytest.raises(ValueError):
        ops.fused_elemwise_add_activation(x, x, ["elementwise_mul", "scale"])


def test_partition_rows_covers_everything():
    for total in (0, 1, 7, 8, 524288 * 8 + 3):
        for world in (1, 2, 3, 8):
            spans = [pdist.partition_rows(total, world, r) for r in range(world)]
            assert spans[0][0] == 0
            for (s0, c0), (s1, _c1) in zip(spans, spans[1:]):
                assert s0 + c0 == s1
            assert spans[-1][0] + spans[-1][1] == total
            counts = [c for _, c in spans]
            assert max(counts) - min(counts) <= 1
