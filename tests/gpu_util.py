"""Helpers for the -m gpu parity tests: numpy <-> device, tolerances from the north star.

fp32 reductions: relative 1e-5, taken against sum(|x|) over the reduced slice (the
conditioning-aware form of "rel 1e-5": SURVEY.md §7.2) plus exact equality for
max/min; bf16/fp16: relative 1e-2; integers, bools and indices: bit-exact."""
import numpy as np
import torch

FP32_RTOL = 1e-5
FP64_RTOL = 1e-12
BF16_RTOL = 1e-2
FP16_RTOL = 1e-2


def to_dev(x):
    return torch.from_numpy(np.array(x, order="C")).cuda()   # keeps 0-D arrays 0-D


def to_np(t):
    if t.dtype == torch.bfloat16:
        return t.float().cpu().numpy()
    return t.cpu().numpy()


def bf16_round(x):
    """float32 -> nearest bf16 (as float32 values) -- how the test inputs are made."""
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).bfloat16().float().numpy()


def assert_reduce_close(got, want, x, axis, op, rtol):
    got = np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, (got.shape, want.shape)
    if op in ("max", "min", "amax", "amin") or not np.issubdtype(want.dtype, np.floating):
        assert np.array_equal(got, want), op
        return
    ax = None if axis is None else tuple(np.atleast_1d(axis).tolist())
    if x.ndim == 0:
        ax = None
    if op == "prod":
        scale = np.abs(want).astype(np.float64)
    else:
        scale = np.sum(np.abs(x.astype(np.float64)), axis=ax if ax != () else None,
                       keepdims=False).reshape(want.shape)
        if op == "mean":
            n = x.size // max(want.size, 1)
            scale = scale / max(n, 1)
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    # The reference's CPU kernel accumulates a partial reduction one element after another
    # in the tensor's own precision (Eigen GenericDimReducer), so for long reductions ITS
    # result is further than rtol from the exact value.  The bound is widened by exactly
    # that distance: |gpu - ref| <= rtol*sum|x| + |ref - exact|.
    ref_err = 0.0
    if op in ("sum", "mean"):
        exact = np.sum(x.astype(np.float64), axis=ax if ax != () else None).reshape(want.shape)
        if op == "mean":
            exact = exact / max(x.size // max(want.size, 1), 1)
        ref_err = np.abs(want.astype(np.float64) - exact)
        assert np.all(np.abs(got.astype(np.float64) - exact) <= rtol * scale + 1e-30), \
            (op, "gpu vs exact", float(np.max(np.abs(got - exact) / (scale + 1e-30))))
    bound = rtol * scale + ref_err + 1e-30
    assert np.all(err <= bound), (op, float(np.max(err / (scale + 1e-30))))
