"""Pins the C restatement (oracle/phi_cpu_oracle.c) to the reference.

 - against tests/golden/phi_cpu_golden.npz (outputs of the reference's own CPU kernels, recorded by
   tests/golden/make_golden.py) — always runs;
 - against oracle/_ref/libphi_cpu_ref.so live, on fresh random inputs — when that library exists.
Integer, max/min, argmax and elementwise results must be bit-identical; float sum/mean are
bit-identical too (the Eigen accumulation order is restated); exp-based and prod results to 1e-6.
"""
import os

import numpy as np
import pytest

import oracle_lib as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_cpu_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLDEN))


def _cases(gold, prefix):
    return sorted({k.split("__")[0] for k in gold if k.startswith(prefix) and "__in0" in k})


def _axis_of(name, table):
    for n, shape, axis, *rest in table:
        if n == name:
            return axis, (rest[0] if rest else False)
    raise KeyError(name)


from golden.make_golden import (BCAST_CASES, GRAD_LN_CASES, GRAD_SOFTMAX_CASES, REDUCE_CASES,  # noqa: E402
                                SIBLING_CASES, SIBLING_OPS)


def test_doc_known_answers(port, gold):
    # python/paddle/tensor/math.py:1530-1585
    x = gold["doc_sum__in0"]
    np.testing.assert_allclose(port.reduce("sum", x), 3.5, rtol=1e-6)
    np.testing.assert_allclose(port.reduce("sum", x, 0), [0.3, 0.5, 1.1, 1.6], rtol=1e-6)
    np.testing.assert_allclose(port.reduce("sum", x, -1), [1.9, 1.6], rtol=1e-6)
    assert port.reduce("sum", x, -1, True).shape == (2, 1)
    for k in ("doc_sum_all", "doc_sum_0", "doc_sum_m1", "doc_sum_m1_keep"):
        ax = {"doc_sum_all": None, "doc_sum_0": 0}.get(k, -1)
        assert np.array_equal(port.reduce("sum", x, ax, k.endswith("keep")), gold[k + "__out"])
    # int32 sum -> int64 by InferMeta: the kernel sees an int64 tensor (cpu/reduce.h:56-65)
    y = np.array([[1, 2], [3, 4]], np.int32).astype(np.int64)
    assert port.reduce("sum", y).item() == 10


def test_reduce_golden(port, gold):
    n = 0
    for name, shape, axis, keep in REDUCE_CASES:
        for dt in ("float32", "float64", "int32", "int64"):
            x = gold[f"red_{name}_{dt}__in0"]
            for op in ("sum", "mean", "max", "min", "prod"):
                want = gold[f"red_{name}_{dt}_{op}__out"]
                got = port.reduce(op, x, axis, keep)
                assert got.shape == want.shape
                if dt.startswith("int") or op in ("max", "min", "sum", "mean"):
                    assert np.array_equal(got, want), (name, dt, op)
                else:
                    np.testing.assert_allclose(got, want, rtol=1e-6)
                n += 1
    assert n == len(REDUCE_CASES) * 4 * 5
    x = gold["red_wrap_int64__in0"]
    assert np.array_equal(port.reduce("prod", x, 1), gold["red_wrap_int64_prod__out"])
    assert np.array_equal(port.reduce("sum", x), gold["red_wrap_int64_sum__out"])


def test_elementwise_golden(port, gold):
    for name, xs, ys in BCAST_CASES:
        for dt in ("float32", "int32", "int64", "float64", "int16", "uint8", "int8", "bool", "bfloat16"):
            a, b = gold[f"ew_{name}_{dt}__in0"], gold[f"ew_{name}_{dt}__in1"]
            code = O.BFLOAT16 if dt == "bfloat16" else None
            for op in ("add", "multiply"):
                got = port.elementwise(op, a, b, code)
                want = gold[f"ew_{name}_{dt}_{op}__out"]
                assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), (name, dt, op)
    assert np.array_equal(port.f32_to_bf16(gold["bf16_round__in0"]), gold["bf16_round__out"])


def test_sibling_elementwise_golden(port, gold):
    n = 0
    for name, xs, ys in SIBLING_CASES:
        for dt, ops in SIBLING_OPS.items():
            a, b = gold[f"ew2_{name}_{dt}__in0"], gold[f"ew2_{name}_{dt}__in1"]
            code = O.BFLOAT16 if dt == "bfloat16" else None
            for op in ops:
                got = port.elementwise(op, a, b, code)
                want = gold[f"ew2_{name}_{dt}_{op}__out"]
                assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), (name, dt, op)
                n += 1
    assert n == len(SIBLING_CASES) * 22
    with pytest.raises(ValueError):  # integer division by zero is an error on the CPU (elementwise_functor.h:98-115)
        port.elementwise("divide", np.ones(3, np.int32), np.zeros(3, np.int32))


def test_argmax_golden(port, gold):
    table = {"a3_0": 0, "a3_1": 1, "a3_m1": -1, "a_flat": None, "a1": 0, "a2_wide": 1}
    for name, axis in table.items():
        for dt in ("float32", "float64", "int32", "int64", "int16", "uint8"):
            x = gold[f"arg_{name}_{dt}__in0"]
            assert np.array_equal(port.argmax(x, axis, False, np.int64), gold[f"arg_{name}_{dt}_i64__out"])
            got = port.argmax(x, axis, True, np.int32)
            want = gold[f"arg_{name}_{dt}_i32_keep__out"]
            assert got.dtype == np.int32 and got.shape == want.shape and np.array_equal(got, want)


def test_softmax_golden(port, gold):
    table = {"s2": -1, "s4_m1": -1, "s4_0": 0, "s4_1": 1, "s4_2": 2, "s5": 3, "s_2048": -1, "s_odd": -1, "s_clip": -1}
    for name, axis in table.items():
        for dt in ("float32", "float64"):
            x = gold[f"sm_{name}_{dt}__in0"]
            got = port.softmax(x, axis)
            np.testing.assert_allclose(got, gold[f"sm_{name}_{dt}__out"], rtol=2e-6 if dt == "float32" else 1e-14)
    # the -64 clip is visible: exp(-64)/sum, not exp(-1e4)/sum == 0
    y = port.softmax(gold["sm_s_clip_float32__in0"], -1)
    assert y[0, 3] > 0


def test_layer_norm_golden(port, gold):
    table = {"l4": 1, "l4_3": 3, "l2_8192": 1, "l2_odd": 1, "l3": 2}
    for name, bna in table.items():
        for dt in ("float32", "float64"):
            key = f"ln_{name}_{dt}"
            x, w, b = gold[key + "__in0"], gold[key + "__in1"], gold[key + "__in2"]
            y, m, v = port.layer_norm(x, bna, w, b, 1e-5)
            assert np.array_equal(m, gold[key + "_wb__out1"])
            assert np.array_equal(v, gold[key + "_wb__out2"])
            np.testing.assert_allclose(y, gold[key + "_wb__out0"], rtol=1e-6, atol=1e-7)
            y0, _, _ = port.layer_norm(x, bna, None, None, 1e-5)
            np.testing.assert_allclose(y0, gold[key + "_none__out0"], rtol=1e-6, atol=1e-7)


def test_grad_kernels_golden(port, gold):
    for name, shape, axis in GRAD_SOFTMAX_CASES:
        for dt in ("float32", "float64"):
            key = f"smg_{name}_{dt}"
            got = port.softmax_grad(gold[key + "__in0"], gold[key + "__in1"], axis)
            np.testing.assert_allclose(got, gold[key + "__out"], rtol=1e-6 if dt == "float32" else 1e-13, atol=1e-30)
    for name, shape, bna in GRAD_LN_CASES:
        for dt in ("float32", "float64"):
            key = f"lng_{name}_{dt}"
            x, w, m, v, dy = (gold[key + f"__in{i}"] for i in range(5))
            dx, ds, db = port.layer_norm_grad(x, bna, w, m, v, dy, 1e-5)
            assert np.array_equal(dx, gold[key + "_w__out0"]) and np.array_equal(ds, gold[key + "_w__out1"])
            assert np.array_equal(db, gold[key + "_w__out2"])
            dx, _, db = port.layer_norm_grad(x, bna, None, m, v, dy, 1e-5)
            assert np.array_equal(dx, gold[key + "_now__out0"]) and np.array_equal(db, gold[key + "_now__out2"])


# ---- live comparison with the compiled reference ----------------------------------------------
def test_live_reduce(port, ref):
    rng = np.random.default_rng(7)
    x = rng.random((4096, 4096), dtype=np.float32)  # BASELINE config 1
    assert np.array_equal(port.reduce("sum", x, -1), ref.reduce("sum", x, -1))
    assert np.array_equal(port.reduce("sum", x), ref.reduce("sum", x))
    assert np.array_equal(port.reduce("mean", x, 0), ref.reduce("mean", x, 0))
    for n in (1, 7, 8, 9, 1000, 8191, 8192, 8193, 100003):
        v = rng.standard_normal(n).astype(np.float32)
        for op in ("sum", "mean", "max", "min"):
            assert np.array_equal(port.reduce(op, v), ref.reduce(op, v)), (n, op)
    xi = rng.integers(-2**62, 2**62, (64, 1031), dtype=np.int64)
    for op in ("sum", "max", "min", "prod", "mean"):
        for ax in (0, 1, None):
            assert np.array_equal(port.reduce(op, xi, ax), ref.reduce(op, xi, ax))


def test_live_elementwise_and_halves(port, ref):
    rng = np.random.default_rng(8)
    f = rng.standard_normal(100000).astype(np.float32) * 100
    assert np.array_equal(port.f32_to_bf16(f), ref.f32_to_bf16(f))
    a, b = port.f32_to_bf16(f[:4096 * 4].reshape(4, 4096)), port.f32_to_bf16(f[4096 * 4:4096 * 5])
    for op in ("add", "multiply"):
        assert np.array_equal(port.elementwise(op, a, b, O.BFLOAT16), ref.elementwise(op, a, b, O.BFLOAT16))
    # fp16: the port rounds to nearest even (== numpy, == the GPU's __float2half_rn, == an F16C
    # build).  Paddle's default CPU build has no -mf16c and falls back to a TRUNCATING software
    # routine (phi/common/float16.h:121-138), so the compiled reference may sit 1 ulp below.
    h = port.f32_to_fp16(f)
    assert np.array_equal(h, f.astype(np.float16).view(np.uint16))
    hr = ref.f32_to_fp16(f)
    assert np.max(np.abs(h.astype(np.int32) - hr.astype(np.int32))) <= 1


def test_live_softmax_layernorm_argmax(port, ref):
    rng = np.random.default_rng(9)
    s = (rng.standard_normal((32, 2048)) * 2).astype(np.float32)
    np.testing.assert_allclose(port.softmax(s), ref.softmax(s), rtol=2e-6)
    x = (rng.standard_normal((16, 8192)) * 3 + 1).astype(np.float32)
    w, b = rng.random(8192, dtype=np.float32), rng.random(8192, dtype=np.float32)
    for p_, r_ in zip(port.layer_norm(x, 1, w, b), ref.layer_norm(x, 1, w, b)):
        np.testing.assert_allclose(p_, r_, rtol=1e-6, atol=1e-7)
    t = rng.integers(0, 9, (7, 300, 5)).astype(np.float32)
    for ax in (0, 1, 2, None):
        assert np.array_equal(port.argmax(t, ax), ref.argmax(t, ax))


def test_live_argmax_nan_and_infinities(port, ref):
    """Eigen's accumulator starts as (0, lowest()): NaNs never win wherever they stand, and a row that never
    beats lowest() (all -inf / NaN / lowest) answers 0."""
    rng = np.random.default_rng(10)
    for dt in (np.float32, np.float64):
        x = rng.choice(np.array([-np.inf, -3.0, -0.0, 0.0, 0.5, 2.0, np.nan], dt), size=(40, 37))
        x[3] = np.nan
        x[4] = -np.inf
        x[5] = np.finfo(dt).min
        x[6, 0] = np.nan
        for ax in (0, 1, None):
            assert np.array_equal(port.argmax(x, ax), ref.argmax(x, ax))


def test_argmin_restatement_equals_compiled_reference(port, ref):
    """argmin (added in round 2) has no golden file: the restatement is pinned to the compiled reference live, over the
    dtypes the CPU kernel registers, every axis and the flattened form, with ties, NaN, +-inf and all-equal rows."""
    rng = np.random.default_rng(21)
    for dt in (np.float32, np.float64, np.int32, np.int64, np.int16, np.uint8):
        x = rng.integers(0, 12, (6, 31, 5)).astype(dt)
        if np.issubdtype(dt, np.floating):
            x[0, 3, 1] = np.nan
            x[1, :, 2] = np.inf
            x[2, 0, 0] = -np.inf
            x[3, :, 3] = np.nan
        for axis in (None, 0, 1, 2, -1):
            for keep in (False, True):
                a, b = port.argmin(x, axis, keep), ref.argmin(x, axis, keep)
                assert a.shape == b.shape and np.array_equal(a, b), (dt, axis, keep)
        assert np.array_equal(port.argmin(x, 1, dtype=np.int32), ref.argmin(x, 1, dtype=np.int32))


def test_select_grad_restatement_equals_compiled_reference(port, ref):
    """max_grad / min_grad / prod_grad (round 2) have no golden file: the restatement is pinned to the compiled reference
    live.  max / min: the CPU kernel compares exactly, so the restatement runs with tol=False here and, on inputs whose
    distinct values are further than 1e-8 apart, with the GPU rule (tol=True) as well — the same bits.  prod: bit-exact for
    float32 / float64 (three correctly rounded steps), integers away from x == 0."""
    rng = np.random.default_rng(33)
    for dt in (np.float32, np.float64, np.int32, np.int64):
        for shape, axis in (((6, 31, 5), None), ((6, 31, 5), 1), ((6, 31, 5), (0, 2)), ((6, 31, 5), -1), ((17,), 0), ((4, 1, 9), 1)):
            x = rng.integers(-3, 4, shape).astype(dt)  # many ties
            for keep in (False, True):
                axes = None if axis is None else (list(axis) if isinstance(axis, tuple) else [axis])
                for op, fwd in (("max", np.max), ("min", np.min)):
                    out = fwd(x, axis=axis, keepdims=keep)
                    g = rng.integers(-5, 6, np.shape(out)).astype(dt)
                    if np.issubdtype(dt, np.floating) and g.size:
                        g.flat[0] = np.inf  # inf * 0 = NaN at the non-maximal positions
                    want = ref.select_grad(op, x, out, g, axes, keep)
                    assert np.array_equal(port.select_grad(op, x, out, g, axes, keep, tol=False), want, equal_nan=True), (dt, shape, axis, op)
                    assert np.array_equal(port.select_grad(op, x, out, g, axes, keep, tol=True), want, equal_nan=True), (dt, shape, axis, op)
                xp = x.copy()
                xp[xp == 0] = 2
                if np.issubdtype(dt, np.floating):
                    xp = (xp * rng.uniform(0.5, 1.5, shape)).astype(dt)
                out = np.prod(xp, axis=axis, keepdims=keep).astype(dt)
                g = rng.standard_normal(np.shape(out)).astype(dt) if np.issubdtype(dt, np.floating) else rng.integers(-5, 6, np.shape(out)).astype(dt)
                want = ref.select_grad("prod", xp, out, g, axes, keep)
                assert np.array_equal(port.select_grad("prod", xp, out, g, axes, keep), want), (dt, shape, axis, "prod")
    # NaN and infinities in x: NaN equals nothing, inf == inf
    x = np.array([[1.0, np.inf, np.nan, np.inf], [np.nan, -np.inf, 2.0, 2.0]], np.float32)
    out = np.array([np.inf, 2.0], np.float32)
    g = np.array([3.0, -1.0], np.float32)
    want = ref.select_grad("max", x, out, g, [1], False)
    assert np.array_equal(want, np.array([[0, 3, 0, 3], [0, 0, -1, -1]], np.float32))
    for tol in (False, True):
        assert np.array_equal(port.select_grad("max", x, out, g, [1], False, tol=tol), want)


def test_fused_layer_norm_helper_agrees_with_the_reference_avx512_kernel(port, ref):
    """fused_bias_residual_layernorm: the test oracle is the reference's layer_norm CPU kernel applied to a sum formed the way
    the GPU helper forms it (oracle_lib.fused_bias_dropout_residual_layer_norm, alpha=...).  The reference also ships a CPU
    kernel of this very op (AVX-512, float32): the two must agree to rounding — it adds the residual before the bias and
    uses E[d^2] - mean^2 — which pins the formula (what is summed, what alpha multiplies, biased variance, where epsilon
    goes, rsqrt) to the reference's own code."""
    if not ref.have_avx512():
        pytest.skip("the reference's CPU kernel of this op needs AVX-512")
    rng = np.random.default_rng(44)
    for rows, cols, alpha, has_bias in ((37, 1000, 0.5, True), (8, 4096, 1.0, False), (64, 70, 2.0, True), (3, 16, 0.25, True)):
        x = (rng.standard_normal((rows, cols)) * 2).astype(np.float32)
        res = (rng.standard_normal((rows, cols)) * 2 + 0.5).astype(np.float32)
        bias = rng.uniform(-1, 1, cols).astype(np.float32) if has_bias else None
        w = rng.uniform(0.5, 1.5, cols).astype(np.float32)
        b = rng.uniform(-0.5, 0.5, cols).astype(np.float32)
        wy, wd, wm, wv = ref.fused_bias_dropout_residual_layer_norm(x, res, bias, w, b, 1e-5, 1.0, alpha=alpha)
        out, ro, m, rstd = ref.fused_layer_norm_cpu(x, bias, res, w, b, 1e-5, alpha)
        np.testing.assert_allclose(ro, wd, rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(m, wm, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(rstd, 1.0 / np.sqrt(wv + 1e-5), rtol=2e-5)
        np.testing.assert_allclose(out, wy, rtol=1e-4, atol=1e-4)
        # the sum-only form and the form without a residual
        s, ro2, _, _ = ref.fused_layer_norm_cpu(x, bias, res, None, None, 1e-5, alpha)
        assert ro2 is None
        np.testing.assert_allclose(s, ref.fused_bias_dropout_residual_layer_norm(x, res, bias, None, None, 1e-5, 1.0, alpha=alpha,
                                                                                  sum_only=True), rtol=1e-6, atol=1e-6)
        out0, _, m0, r0 = ref.fused_layer_norm_cpu(x, None, None, w, b, 1e-5, 1.0)
        y0, mm0, vv0 = ref.layer_norm(x, 1, w, b, 1e-5)
        np.testing.assert_allclose(out0, y0, rtol=1e-4, atol=1e-4)
        np.testing.assert_allclose(r0, 1.0 / np.sqrt(vv0 + 1e-5), rtol=2e-5)


def test_log_softmax_grad_restatement_equals_compiled_reference(port, ref):
    """log_softmax_grad (round 2): the restatement against the compiled reference CPU kernel (Eigen expression
    dy - exp(y) * sum(dy)); the reducer's summation order moves the last bits, so 1e-6 of the row scale."""
    rng = np.random.default_rng(55)
    for dt in (np.float32, np.float64):
        for shape, axis in (((7, 33), -1), ((7, 33), 0), ((4, 9, 5), 1), ((2, 3, 4, 5), 2), ((64, 1000), 1), ((5,), 0)):
            y = (rng.standard_normal(shape) - 2.0).astype(dt)  # any tensor is a valid "out" for the kernel
            g = rng.standard_normal(shape).astype(dt)
            a, b = port.log_softmax_grad(y, g, axis), ref.log_softmax_grad(y, g, axis)
            tol = 1e-6 if dt == np.float32 else 1e-13
            np.testing.assert_allclose(a, b, rtol=tol, atol=tol * float(np.abs(g).sum(axis=axis).max() + 1))
    z = np.zeros((), np.float32)
    assert port.log_softmax_grad(z, z + 3, 0) == 0 and ref.log_softmax_grad(z, z + 3, 0) == 0  # 0-D: gradient 0


def test_cross_entropy_grad_restatement_equals_compiled_reference(port, ref):
    """cross_entropy_with_softmax_grad (round 2, use_softmax = True): the restatement is bit-identical to the compiled
    reference CPU kernel — soft labels g * (p - y); hard labels p * g with g subtracted at the label's class, a position
    whose label is ignore_index all zero; class axis last and in the middle."""
    rng = np.random.default_rng(66)
    for dt in (np.float32, np.float64):
        for shape, axis in (((7, 33), -1), ((64, 1000), 1), ((4, 9, 5), 1), ((3, 5, 4, 6), 2), ((5, 1), -1)):
            ax = axis % len(shape)
            lshape = tuple(1 if i == ax else d for i, d in enumerate(shape))
            x = (rng.standard_normal(shape) * 2).astype(dt)
            sm = port.softmax(x, axis)
            g = rng.standard_normal(lshape).astype(dt)
            lbl = rng.integers(0, shape[ax], lshape).astype(np.int64)
            lbl.reshape(-1)[::3] = -100  # ignored positions
            a, b = port.cross_entropy_with_softmax_grad(lbl, sm, g, False, -100, axis), ref.cross_entropy_with_softmax_grad(lbl, sm, g, False, -100, axis)
            assert np.array_equal(a, b), (dt, shape, axis, "hard")
            y = rng.dirichlet(np.ones(shape[ax]), size=int(np.prod(lshape))).astype(dt)
            y = np.moveaxis(y.reshape(tuple(d for i, d in enumerate(shape) if i != ax) + (shape[ax],)), -1, ax).copy()
            a, b = port.cross_entropy_with_softmax_grad(y, sm, g, True, -100, axis), ref.cross_entropy_with_softmax_grad(y, sm, g, True, -100, axis)
            assert np.array_equal(a, b), (dt, shape, axis, "soft")
