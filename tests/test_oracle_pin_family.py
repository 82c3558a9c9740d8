"""Pins oracle/phi_cpu_oracle_family.c (compare / logical / bitwise / not / scale / cast / where) to the reference:
against tests/golden/phi_family_golden.npz (outputs of the reference's own CPU kernels, recorded by
tests/golden/make_golden_family.py) — always; and live against oracle/_ref/libphi_cpu_ref.so when it exists.
Everything is bit-identical except float32/float64 scale (the reference build may or may not fuse the
multiply-add: one unit in the last place)."""
import os

import numpy as np
import pytest

import oracle_lib as O
from golden.make_golden_family import BCAST, COMPARE_OPS, INT_TYPES, SCALE_ARGS, cast_input, draw

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_family_golden.npz")
HALVES = {"bfloat16": O.BFLOAT16, "float16": O.FLOAT16}


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLDEN))


def test_compare_logical_bitwise_golden(port, gold):
    n = 0
    for name, _, _ in BCAST:
        for dt in INT_TYPES + ["float32", "float64", "bfloat16", "float16"]:
            x, y = gold[f"fam_{name}_{dt}__in0"], gold[f"fam_{name}_{dt}__in1"]
            code = HALVES.get(dt)
            for op in COMPARE_OPS:
                assert np.array_equal(port.compare(op, x, y, code=code), gold[f"cmp_{op}_{name}_{dt}__out"]), (op, name, dt)
                n += 1
            if dt in HALVES:
                continue
            if dt != "uint8":
                for op in ("and", "or", "xor"):
                    assert np.array_equal(port.logical(op, x, y), gold[f"log_{op}_{name}_{dt}__out"]), (op, name, dt)
            if dt in INT_TYPES:
                for op in ("and", "or", "xor"):
                    assert np.array_equal(port.bitwise(op, x, y), gold[f"bit_{op}_{name}_{dt}__out"]), (op, name, dt)
    assert n == len(BCAST) * 10 * 6


def test_unary_golden(port, gold):
    for dt in INT_TYPES + ["float32", "float64"]:
        x = gold[f"un_{dt}__in0"]
        if dt != "uint8":
            assert np.array_equal(port.logical_not(x), gold[f"lnot_{dt}__out"])
        if dt in INT_TYPES:
            assert np.array_equal(port.bitwise_not(x), gold[f"bnot_{dt}__out"])


def test_scale_golden(port, gold):
    for dt in ["uint8", "int8", "int16", "int32", "int64", "float32", "float64", "bool", "bfloat16"]:
        x = gold[f"scale_{dt}__in0"]
        for k, (s, b, after) in enumerate(SCALE_ARGS):
            got, want = port.scale(x, s, b, after, code=HALVES.get(dt)), gold[f"scale_{dt}_{k}__out"]
            if dt in ("float32", "float64"):
                np.testing.assert_allclose(got, want, rtol=2 * np.finfo(dt).eps, atol=np.finfo(dt).eps)
            else:
                assert np.array_equal(got, want), (dt, k)


def test_cast_golden(port, gold):
    for sc in range(10):
        x = gold[f"cast_{sc}__in0"]
        for dc in range(10):
            if dc == O.FLOAT16 and sc != O.FLOAT16:
                continue
            assert np.array_equal(port.cast(x, dc, code=sc), gold[f"cast_{sc}_{dc}__out"]), (sc, dc)


def test_where_golden(port, gold):
    for dt in ("int32", "int64", "float32", "float64"):
        c, x, y = gold[f"where_{dt}__in0"], gold[f"where_{dt}__in1"], gold[f"where_{dt}__in2"]
        assert np.array_equal(port.where(c, x, y), gold[f"where_{dt}__out"])
    # other element sizes: a bit select
    rng = np.random.default_rng(3)
    for dt in (np.uint8, np.int16):
        c, x, y = rng.random(100) < 0.5, rng.integers(0, 99, 100).astype(dt), rng.integers(0, 99, 100).astype(dt)
        assert np.array_equal(port.where(c, x, y), np.where(c, x, y))


def test_live_family(port, ref):
    rng = np.random.default_rng(77)
    for dt in INT_TYPES + ["float32", "float64"]:
        for xs, ys in (((33, 1, 65), (7, 65)), ((1000,), (1000,)), ((5, 6), ())):
            x, y = draw(rng, dt, xs), draw(rng, dt, ys)
            for op in COMPARE_OPS:
                assert np.array_equal(port.compare(op, x, y), ref.compare(op, x, y))
                assert np.array_equal(port.compare(op, y, x), ref.compare(op, y, x))
            if dt != "uint8":
                for op in ("and", "or", "xor"):
                    assert np.array_equal(port.logical(op, x, y), ref.logical(op, x, y))
                assert np.array_equal(port.logical_not(x), ref.logical_not(x))
            if dt in INT_TYPES:
                for op in ("and", "or", "xor"):
                    assert np.array_equal(port.bitwise(op, x, y), ref.bitwise(op, x, y))
                assert np.array_equal(port.bitwise_not(x), ref.bitwise_not(x))
    for sc in range(10):
        x = cast_input(rng, port, sc)
        for dc in range(10):
            if dc == O.FLOAT16 and sc != O.FLOAT16:
                continue
            assert np.array_equal(port.cast(x, dc, code=sc), ref.cast(x, dc, code=sc)), (sc, dc)
    xb = port.f32_to_bf16(rng.standard_normal(4096).astype(np.float32))
    for s, b, after in SCALE_ARGS:
        assert np.array_equal(port.scale(xb, s, b, after, code=O.BFLOAT16), ref.scale(xb, s, b, after, code=O.BFLOAT16))
        xi = rng.integers(-2**40, 2**40, 999)
        assert np.array_equal(port.scale(xi, s, b, after), ref.scale(xi, s, b, after))


def test_elementwise_grad_golden_and_live(port, gold):
    from golden.make_golden_family import GRAD_CASES
    for name, _, _ in GRAD_CASES:
        for dt in ("float32", "float64", "int32", "int64", "bfloat16"):
            x, y, d = (gold[f"grad_{name}_{dt}__in{i}"] for i in range(3))
            for op in ("add", "multiply"):
                if dt == "bfloat16" and op == "add":
                    continue  # no bf16 add_grad kernel on CPU
                dx, dy = port.elementwise_grad(op, x, y, d, code=HALVES.get(dt))
                # every case bit-for-bit (both gradients wanted: the fused multiply-add loops, see the oracle's header)
                assert np.array_equal(dx, gold[f"grad_{op}_{name}_{dt}__out0"]), (op, name, dt)
                assert np.array_equal(dy, gold[f"grad_{op}_{name}_{dt}__out1"]), (op, name, dt)


def test_elementwise_grad_live_single_outputs(port, ref):
    """Only one gradient wanted: the reference build compiles one loop unfused (a few ulp), the rest bit-exact."""
    from golden.make_golden_family import GRAD_CASES
    rng = np.random.default_rng(12)
    for _, xs, ys in GRAD_CASES:
        os_ = np.broadcast_shapes(xs, ys)
        for dt in (np.float32, np.float64, np.int64):
            if dt == np.int64:
                x, y, d = (rng.integers(-2**31, 2**31, s_) for s_ in (xs, ys, os_))
            else:
                x, y, d = (rng.standard_normal(s_).astype(dt) for s_ in (xs, ys, os_))
            for op in ("add", "multiply"):
                for wx, wy in ((True, False), (False, True)):
                    got, want = port.elementwise_grad(op, x, y, d, wx, wy), ref.elementwise_grad(op, x, y, d, wx, wy)
                    for a, b in zip(got, want):
                        if a is None:
                            assert b is None
                        elif dt == np.int64 or op == "add":
                            assert np.array_equal(a, b)
                        else:
                            np.testing.assert_allclose(a, b, rtol=64 * np.finfo(dt).eps, atol=64 * np.finfo(dt).eps)


def test_reduce_grad_golden_and_live(port, gold):
    from golden.make_golden_family import REDUCE_GRAD_CASES
    for name, shape, axis, keep in REDUCE_GRAD_CASES:
        for dt in ("float32", "float64", "int32", "int64"):
            d = gold[f"rgrad_{name}_{dt}__in0"]
            for op in ("sum", "mean"):
                assert np.array_equal(port.reduce_grad(op, shape, d, axis, keep), gold[f"rgrad_{op}_{name}_{dt}__out"]), (op, name, dt)


def test_log_softmax_golden(port, gold):
    from golden.make_golden_family import LOG_SOFTMAX_CASES
    for name, _, axis in LOG_SOFTMAX_CASES:
        for dt in ("float32", "float64"):
            x = gold[f"lsm_{name}_{dt}__in0"]
            np.testing.assert_allclose(port.log_softmax(x, axis), gold[f"lsm_{name}_{dt}__out"],
                                       rtol=2e-6 if dt == "float32" else 1e-12, atol=2e-6 if dt == "float32" else 1e-12)
    assert port.log_softmax(np.array(3.0, np.float32)).item() == 0.0


def test_cross_entropy_golden(port, gold):
    from golden.make_golden_family import CE_CASES
    for name, _, axis in CE_CASES:
        for dt in ("float32", "float64"):
            x, hard, soft = (gold[f"ce_{name}_{dt}__in{i}"] for i in range(3))
            tol = dict(rtol=2e-6, atol=2e-6) if dt == "float32" else dict(rtol=1e-12, atol=1e-12)
            for kind, lbl in (("hard", hard), ("soft", soft)):
                sm, loss = port.cross_entropy_with_softmax(x, lbl, kind == "soft", True, -100, axis)
                np.testing.assert_allclose(sm, gold[f"ce_{kind}_{name}_{dt}__out0"], **tol)
                np.testing.assert_allclose(loss, gold[f"ce_{kind}_{name}_{dt}__out1"], **tol)
                assert loss.shape == gold[f"ce_{kind}_{name}_{dt}__out1"].shape
            assert port.cross_entropy_with_softmax(x, hard, False, True, -100, axis)[1].flat[0] == 0  # ignore_index


def test_bitwise_shift_golden(port, gold):
    from golden.make_golden_family import SHIFT_CASES
    for name, _, _ in SHIFT_CASES:
        for dt in ("uint8", "int8", "int16", "int32", "int64"):
            x, y = gold[f"sh_{name}_{dt}__in0"], gold[f"sh_{name}_{dt}__in1"]
            for right in (0, 1):
                for ar in (0, 1):
                    assert np.array_equal(port.bitwise_shift(right, x, y, bool(ar)), gold[f"sh_{name}_{dt}_{right}{ar}__out"]), (name, dt, right, ar)


def test_rms_norm_restatement_matches_the_reference_tests_formula(port):
    """rms_norm has no CPU kernel in the reference (parity unpinned against oracle/_ref); the restatement is held to the
    formula the reference's own test uses, test/legacy_test/test_rms_norm_op.py:37-47,74-79, written out in numpy."""
    rng = np.random.default_rng(5)
    x, res = rng.standard_normal((6, 7, 40)), rng.standard_normal((6, 7, 40))
    g, b, bias = rng.standard_normal(40), rng.standard_normal(40), rng.standard_normal(40)
    out, _, inv = port.rms_norm(x, g, b, 1e-5, 2)
    var = (x ** 2).mean(-1, keepdims=True)
    np.testing.assert_allclose(out, 1.0 / np.sqrt(var + 1e-5) * x * g + b, rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(inv, 1.0 / np.sqrt(var[..., 0] + 1e-5), rtol=1e-12, atol=1e-14)
    out, ro, _ = port.rms_norm(x, g, b, 1e-5, 2, bias=bias, residual=res)
    v = x + res + bias
    np.testing.assert_allclose(ro, v, rtol=1e-15)
    np.testing.assert_allclose(out, 1.0 / np.sqrt((v ** 2).mean(-1, keepdims=True) + 1e-5) * v * g + b, rtol=1e-12, atol=1e-14)
    out2, _, _ = port.rms_norm(x.reshape(6, 280), np.tile(g, 7), None, 1e-5, 1)
    assert out2.shape == (6, 280)
