"""Broadcast / elementwise / fused add+activation parity through the C ABI."""
import numpy as np
import pytest
import torch

import cases
import gpu_util as G
import paddle2_b200 as pb
from golden.make_golden import parse_functor_list

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("i", range(len(cases.ADD_CASES)))
@pytest.mark.parametrize("dt", ["float32", "float64", "int32", "int64"])
def test_broadcast_add_replay(oracle, i, dt):
    xs, ys, axis, src = cases.ADD_CASES[i]
    rng = np.random.default_rng(3000 + i)
    if np.issubdtype(np.dtype(dt), np.integer):
        x = rng.integers(-50, 50, size=xs).astype(dt)
        y = rng.integers(-50, 50, size=ys).astype(dt)
    else:
        x = cases.make_input(rng, dt, xs, "signed")
        y = cases.make_input(rng, dt, ys, "signed")
    want = oracle.elementwise("add", [x, y], axis=axis)
    got = G.to_np(pb.add(G.to_dev(x), G.to_dev(y), axis=axis))
    assert got.shape == want.shape, src
    assert np.array_equal(got, want), src   # one IEEE add per element: bit-exact


@pytest.mark.parametrize("name", ["add", "subtract", "multiply", "divide", "maximum", "minimum"])
def test_binary_functors(oracle, name):
    rng = np.random.default_rng(17)
    for xs, ys in [((257, 1031), (257, 1031)), ((64, 3, 500), (3, 1)), ((500, 7), (500, 1)), ((9, 1000), (1000,))]:
        x = (rng.random(xs) * 4 - 2).astype(np.float32)
        y = (rng.random(ys) * 4 + 0.5).astype(np.float32)
        got = G.to_np(getattr(pb, name)(G.to_dev(x), G.to_dev(y)))
        want = oracle.elementwise(name, [x, y])
        np.testing.assert_allclose(got, want, rtol=1e-6 if name == "divide" else 0, atol=0)
    xi = rng.integers(-9, 9, size=(33, 65)).astype(np.int64)
    yi = rng.integers(1, 9, size=(65,)).astype(np.int64)
    if name != "divide":
        assert np.array_equal(G.to_np(getattr(pb, name)(G.to_dev(xi), G.to_dev(yi))),
                              oracle.elementwise(name, [xi, yi]))


def test_fused_add_activation_replay(oracle):
    k = 0
    for (xs, ys, axis, src) in cases.FUSED_SHAPES:
        for fl in cases.FUSED_FUNCTOR_LISTS:
            rng = np.random.default_rng(4000 + k)
            x = cases.make_input(rng, "float32", xs, "signed")
            y = cases.make_input(rng, "float32", ys, "signed")
            binary, act, compound = parse_functor_list(fl)
            want, winter = oracle.elementwise(binary, [x, y], act=act, compound=compound, scale=0.1,
                                              axis=axis, want_intermediate=True)
            for save in (False, True):
                res = pb.fused_elemwise_activation(G.to_dev(x), G.to_dev(y), fl, axis=axis, scale=0.1,
                                                   save_intermediate_out=save)
                out = G.to_np(res[0] if save else res)
                np.testing.assert_allclose(out, want, rtol=1e-5, atol=1e-6, err_msg=f"{src} {fl}")
                if save:
                    inter = G.to_np(res[1])
                    if compound:   # intermediate = act(y), shaped like y
                        np.testing.assert_allclose(
                            inter, oracle.elementwise(binary, [y], act=act, scale=0.1), rtol=1e-5, atol=1e-6)
                    else:
                        np.testing.assert_allclose(inter, winter, rtol=1e-5, atol=1e-6)
            k += 1


def test_relu_add_is_bit_exact_up_to_zero_sign(oracle):
    """C2 in small: [8,64,28,28] + [64,1,1]; relu(x+y) has one add and one select."""
    rng = np.random.default_rng(5)
    x = (rng.random((8, 64, 28, 28)) * 2 - 1).astype(np.float32)
    y = (rng.random((64, 1, 1)) * 2 - 1).astype(np.float32)
    got = G.to_np(pb.fused_elemwise_add_activation(G.to_dev(x), G.to_dev(y), ["relu", "elementwise_add"]))
    want = oracle.elementwise("add", [x, y], act="relu")
    assert np.array_equal(got, want)          # -0.0 == +0.0 under array_equal
    two = G.to_np(pb.relu(pb.add(G.to_dev(x), G.to_dev(y))))
    assert np.array_equal(two, want)


def test_ternary_broadcast(oracle):
    # test/cpp/phi/kernels/test_ternary_broadcast.cu:96-120
    rng = np.random.default_rng(8)
    a = rng.random((1, 2048, 3584)).astype(np.float32)
    b = rng.random((1, 2048, 1)).astype(np.float32)
    c = rng.random((1, 1, 3584)).astype(np.float32)
    got = G.to_np(pb.add_n3(G.to_dev(a), G.to_dev(b), G.to_dev(c)))
    assert np.array_equal(got, (a + b) + c)
    for dt in (torch.bfloat16, torch.float16):
        at, bt, ct = (torch.from_numpy(v).cuda().to(dt) for v in (a, b, c))
        want = ((at.float() + bt.float()).to(dt).float() + ct.float()).to(dt)
        got = pb.add_n3(at, bt, ct)
        # the kernel keeps (a+b) in fp32 before adding c: allow one 16-bit ulp
        np.testing.assert_allclose(got.float().cpu().numpy(), want.float().cpu().numpy(), rtol=1e-2)


def test_bf16_add_rounds_once(oracle):
    rng = np.random.default_rng(4)
    x = torch.from_numpy(rng.standard_normal((257, 1031)).astype(np.float32)).cuda().bfloat16()
    y = torch.from_numpy(rng.standard_normal((1031,)).astype(np.float32)).cuda().bfloat16()
    got = pb.add(x, y)
    want = (x.float() + y.float()).bfloat16()    # phi bfloat16 operator+: add in fp32, round once
    assert torch.equal(got, want)


def test_misaligned_and_odd_rows(oracle):
    rng = np.random.default_rng(21)
    base = rng.random(1000 * 1031 + 3).astype(np.float32)
    xt = G.to_dev(base)[3:].view(1000, 1031)     # data pointer off the 16-byte grid
    y = rng.random((1031,)).astype(np.float32)
    got = G.to_np(pb.add(xt, G.to_dev(y)))
    assert np.array_equal(got, base[3:].reshape(1000, 1031) + y)


def test_errors(oracle):
    with pytest.raises(ValueError, match="Broadcast dimension mismatch"):
        pb.add(torch.ones(2, 3, device="cuda"), torch.ones(2, 4, device="cuda"))
    with pytest.raises(TypeError):
        pb.add(torch.ones(2, device="cuda"), torch.ones(2, device="cuda", dtype=torch.float64))


# ---- mixed-precision add (MultiPrecisionAddKernelImpl, kps/elementwise_kernel.cu:54-99) ---------
def _bits_to_torch(bits, bf16):
    t = torch.from_numpy(bits.view(np.int16).copy())
    return t.view(torch.bfloat16 if bf16 else torch.float16)


def test_mixed_add_golden():
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "phi_ref_golden.npz"))
    x = G.to_dev(g["mpadd_x"])
    for key, bf16 in (("bf16", True), ("fp16", False)):
        y = _bits_to_torch(g[f"mpadd_{key}_bits"], bf16).cuda()
        got = pb.add(x, y)
        assert got.dtype == torch.float32
        assert np.array_equal(G.to_np(got), g[f"mpadd_{key}_out"]), key   # one IEEE add: bit-exact


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("xs,ys", [((4096, 4096), (4096, 4096)), ((257, 1031), (257, 1031)), ((64, 3, 500), (3, 1)),
                                   ((500, 77), (77,)), ((9, 1000), (1,)), ((1, 4100), (33, 1)), ((7, 5, 3), (7, 1, 3))])
def test_mixed_add_shapes(tdt, xs, ys):
    gen = torch.Generator().manual_seed(4)
    x = torch.randn(xs, generator=gen)
    y = torch.randn(ys, generator=gen).to(tdt)
    got = pb.add(x.cuda(), y.cuda())
    want = x + y.float()          # exact: widening is exact, the add is one rounding
    assert got.shape == want.shape
    assert torch.equal(got.cpu(), want)


def test_mixed_add_in_place_accumulate():
    """master_grad += grad: out aliases x"""
    gen = torch.Generator().manual_seed(5)
    acc = torch.randn((1024, 2048), generator=gen).cuda()
    want = acc.clone()
    for _ in range(3):
        gr = torch.randn((1024, 2048), generator=gen).bfloat16().cuda()
        want = want + gr.float()
        r = pb.add(acc, gr, out=acc)
        assert r.data_ptr() == acc.data_ptr()
    assert torch.equal(acc, want)
    with pytest.raises(pb.B200PhiError):
        from paddle2_b200 import _lib as L
        import ctypes as C
        d = L.MixedAddDesc()
        d.y_dtype = L.FLOAT32
        L.check(L.lib().b200phi_mixed_add(C.byref(d), None))
