"""Generates tests/golden/phi_ref_golden.npz from the REFERENCE build of the oracle
(oracle/_ref/libphi_ref.so: the reference's own functors + vendored Eigen, compiled
from /root/reference by oracle/Makefile).  Run in the build container only:

    make -C oracle ref && python tests/golden/make_golden.py

The fixture holds, for every case of tests/cases.py, the input(s) and the output
the reference arithmetic produced, plus the literal answers of the reference's
docstring examples.  The GPU box has no /root/reference; it checks against this file.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402
from oracle import pyoracle as O  # noqa: E402


def main():
    assert O.have_ref(), "build oracle/_ref first (make -C oracle ref)"
    store = {}
    for i, (op, dt, shape, axis, keepdim, _src) in enumerate(cases.REDUCE_CASES):
        rng = np.random.default_rng(123 + i)
        kind = "logical" if op in ("any", "all") else ("prod" if op == "prod" else "uniform")
        x = cases.make_input(rng, dt, shape, kind)
        store[f"reduce_{i}_x"] = x
        store[f"reduce_{i}_y"] = O.reduce(op, x, axis, keepdim, impl="ref")
    for i, (idt, odt, shape, axis, keepdim, _src) in enumerate(cases.SUM_DTYPE_CASES):
        rng = np.random.default_rng(1000 + i)
        x = cases.make_input(rng, idt, shape)
        store[f"sumdt_{i}_x"] = x
        if np.dtype(odt) in (np.dtype(np.float32), np.dtype(np.float64), np.dtype(np.int32),
                             np.dtype(np.int64)):
            store[f"sumdt_{i}_y"] = O.reduce("sum", x, axis, keepdim, out_dtype=odt, impl="ref")
    for i, (op, dt, shape, axis, keepdim, idt, _src) in enumerate(cases.ARG_CASES):
        rng = np.random.default_rng(2000 + i)
        x = (1000 * rng.random(shape)).astype(dt)  # test_arg_min_max_op.py:38
        store[f"arg_{i}_x"] = x
        store[f"arg_{i}_y"] = O.argreduce(op, x, axis, keepdim, np.dtype(idt), impl="ref")
    for i, (xs, ys, axis, _src) in enumerate(cases.ADD_CASES):
        rng = np.random.default_rng(3000 + i)
        x = cases.make_input(rng, "float32", xs, "signed")
        y = cases.make_input(rng, "float32", ys, "signed")
        store[f"add_{i}_x"], store[f"add_{i}_y"] = x, y
        store[f"add_{i}_z"] = O.elementwise("add", [x, y], axis=axis, impl="ref")
    k = 0
    for (xs, ys, axis, _src) in cases.FUSED_SHAPES:
        for fl in cases.FUSED_FUNCTOR_LISTS:
            rng = np.random.default_rng(4000 + k)
            x = cases.make_input(rng, "float32", xs, "signed")
            y = cases.make_input(rng, "float32", ys, "signed")
            binary, act, compound = parse_functor_list(fl)
            out, inter = O.elementwise(binary, [x, y], act=act, compound=compound, scale=0.1,
                                       axis=axis, want_intermediate=True, impl="ref")
            store[f"fused_{k}_x"], store[f"fused_{k}_y"] = x, y
            store[f"fused_{k}_out"], store[f"fused_{k}_inter"] = out, inter
            k += 1
    # ternary broadcast (test/cpp/phi/kernels/test_ternary_broadcast.cu:96-120), scaled down
    rng = np.random.default_rng(5000)
    a = rng.random((1, 64, 112)).astype(np.float32)
    b = rng.random((1, 64, 1)).astype(np.float32)
    c = rng.random((1, 1, 112)).astype(np.float32)
    store["ternary_a"], store["ternary_b"], store["ternary_c"] = a, b, c
    store["ternary_out"] = O.elementwise("add", [a, b, c], impl="ref")
    # row variance (test_variance_layer.py: np.var(ddof=1) oracle; composite stat.py:184-197)
    x = np.random.default_rng(6000).random((64, 1000)).astype(np.float32)
    store["var_x"] = x
    store["var_y"] = O.var_composite(x, axis=-1, unbiased=True, impl="ref")
    store["mean_y"] = O.reduce("mean", x, [-1], impl="ref")
    # functor plug-ins (b200phi_map): the reference's own compare / logical / bitwise / backward
    # functors over the broadcast matrix of test_logical_op.py:45-60, its CPU activation functors
    for i, (xs, ys, _src) in enumerate(cases.MAP_SHAPES):
        for dt in ("float32", "int32", "bool"):
            rng = np.random.default_rng(7000 + i)
            x = cases.map_operand(rng, dt, xs, "ties")
            y = cases.map_operand(rng, dt, ys, "ties")
            store[f"map_{i}_{dt}_x"], store[f"map_{i}_{dt}_y"] = x, y
            for name in cases.COMPARE_OPS + cases.LOGICAL_OPS:
                store[f"map_{i}_{dt}_{name}"] = O.map(name, [x, y], impl="ref")
            if dt != "float32":
                for name in cases.BITWISE_OPS:
                    store[f"map_{i}_{dt}_{name}"] = O.map(name, [x, y], impl="ref")
    rng = np.random.default_rng(7100)
    x = cases.map_operand(rng, "float32", [6, 5, 40], "ties")
    y = cases.map_operand(rng, "float32", [6, 1, 40], "ties")
    g = cases.map_operand(rng, "float32", [6, 5, 40])
    store["mapgrad_x"], store["mapgrad_y"], store["mapgrad_g"] = x, y, g
    for name in ("equal_mul", "max_grad_x", "max_grad_y", "min_grad_x", "min_grad_y", "div_grad_y"):
        store[f"mapgrad_{name}"] = O.map(name, [x, y, g], impl="ref")
    for name, p0, p1, dom in cases.ACTIVATIONS:
        x = cases.map_operand(np.random.default_rng(7200), "float32", [11, 17], dom)
        store[f"act_{name}_x"] = x
        store[f"act_{name}_y"] = O.map(name, [x], p0, p1, impl="ref")
    # reduce-side plug-ins: the Eigen expressions of cpu/p_norm_kernel.cc, squared_l2_norm.h and
    # phi::LogsumexpFunctor itself (oracle/ref_src/ref_norm.cc)
    for i, (shape, axis, porder, keepdim, asvector, _src) in enumerate(cases.PNORM_CASES):
        x = cases.map_operand(np.random.default_rng(8000 + i), "float32", shape)
        x.reshape(-1)[::7] = 0  # zeros for the porder 0 count
        store[f"pnorm_{i}_x"] = x
        store[f"pnorm_{i}_y"] = O.p_norm(x, porder, axis, keepdim, asvector, impl="ref")
    x = cases.map_operand(np.random.default_rng(8100), "float32", [13, 1001])
    store["sql2_x"], store["sql2_y"] = x, O.squared_l2_norm(x, impl="ref")
    for i, (shape, axis, keepdim, _src) in enumerate(cases.LOGSUMEXP_CASES):
        x = np.random.default_rng(8200 + i).uniform(-1, 1, shape).astype(np.float32)  # test_logsumexp.py:71
        store[f"lse_{i}_x"] = x
        axes = [a % max(len(shape), 1) for a in axis]
        if sorted(axes) == list(range(min(axes), min(axes) + len(axes))) if axes else True:
            store[f"lse_{i}_y"] = O.logsumexp(x, axis if axis else None, keepdim, impl="ref")
    # shifts (funcs/bitwise_functors.h:49-206): counts below 0, inside and beyond the bit width
    for dt in cases.SHIFT_DTYPES:
        rng = np.random.default_rng(9100)
        bits = np.dtype(dt).itemsize * 8
        x = cases.map_operand(rng, dt, [7, 40])
        lo = 0 if dt == "uint8" else -3
        y = rng.integers(lo, bits + 4, size=[40]).astype(dt)
        store[f"shift_{dt}_x"], store[f"shift_{dt}_y"] = x, y
        for name in cases.SHIFT_OPS:
            store[f"shift_{dt}_{name}"] = O.map(name, [x, y], impl="ref")
    # the arithmetic binary functors next to add .. minimum in kps/elementwise_kernel.cu:183-293:
    # the reference's own Remainder / FloorDivide / ElementwisePow / CopySign / FMax / FMin /
    # ElementwiseHeaviside functors over a small broadcast matrix
    for name, dtypes in cases.ARITH_OPS:
        for dt in dtypes:
            for j, (xs, ys) in enumerate(cases.ARITH_SHAPES):
                x, y = cases.arith_operands(np.random.default_rng(9500 + j), name, dt, xs, ys)
                store[f"arith_{name}_{dt}_{j}_x"], store[f"arith_{name}_{dt}_{j}_y"] = x, y
                store[f"arith_{name}_{dt}_{j}_out"] = O.map(name, [x, y], impl="ref")
    # mixed-precision add (MultiPrecisionAddFunctor): fp32 + bf16 / fp16 bit patterns
    rng = np.random.default_rng(9000)
    x = cases.map_operand(rng, "float32", [33, 257])
    bits = rng.integers(0, 1 << 16, size=[33, 257]).astype(np.uint16)
    bits[(bits & 0x7F80) == 0x7F80] = 0x3F80  # keep bf16 finite
    hbits = bits.copy()
    hbits[(hbits & 0x7C00) == 0x7C00] = 0x3C00  # keep fp16 finite
    store["mpadd_x"], store["mpadd_bf16_bits"], store["mpadd_fp16_bits"] = x, bits, hbits
    store["mpadd_bf16_out"] = O.mixed_add(x, bits, True, impl="ref")
    store["mpadd_fp16_out"] = O.mixed_add(x, hbits, False, impl="ref")
    np.savez_compressed(os.path.join(HERE, "phi_ref_golden.npz"), **store)
    print("wrote", len(store), "arrays")


def parse_functor_list(fl):
    binary_names = {"elementwise_add": "add", "elementwise_mul": "multiply", "elementwise_sub": "subtract"}
    if fl[0] in binary_names:
        return binary_names[fl[0]], fl[1], 1
    return binary_names[fl[1]], fl[0], 0


if __name__ == "__main__":
    main()
