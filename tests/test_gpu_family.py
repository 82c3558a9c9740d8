"""GPU parity for the rest of the elementwise family: compare / logical / bitwise /
logical_not / bitwise_not / scale / cast / where — CUDA kernels through the Python mirror and the C-ABI, against the
reference's recorded outputs (tests/golden/phi_family_golden.npz) and the CPU oracle.

Bars: bit-exact for everything that is integer, boolean, a comparison, a bit select or a conversion; scale on
float32/float64 to one unit in the last place (fused vs unfused multiply-add); scale on bf16/fp16 to one unit in the
last place of the 16-bit result, or 2^-8 (bf16) / 2^-11 (fp16) of the larger term where the two terms cancel —
the GPU functor computes in float and rounds once, the CPU kernel rounds after each operation.
"""
import os

import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, half_bits_to_f32, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402
from golden.make_golden_family import BCAST, COMPARE_OPS, INT_TYPES, SCALE_ARGS, cast_input, draw  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_family_golden.npz")
HALF = {"bfloat16": "bf16", "float16": "fp16"}
HALF_CODE = {"bfloat16": O.BFLOAT16, "float16": O.FLOAT16}
CODE2TORCH = {O.BOOL: torch.bool, O.UINT8: torch.uint8, O.INT8: torch.int8, O.INT16: torch.int16, O.INT32: torch.int32,
              O.INT64: torch.int64, O.FLOAT16: torch.float16, O.BFLOAT16: torch.bfloat16, O.FLOAT32: torch.float32,
              O.FLOAT64: torch.float64}
CODE_HALF = {O.FLOAT16: "fp16", O.BFLOAT16: "bf16"}


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLDEN))


@pytest.fixture(scope="module", autouse=True)
def _native_code_loaded():
    from paddlepaddle_b200 import _lib
    lib = _lib.load()
    before = lib.pdb200_launch_count()
    yield
    assert lib.pdb200_launch_count() > before, "no pdb200 kernel was launched by this module"


def _logical(op):
    return {"and": paddle.logical_and, "or": paddle.logical_or, "xor": paddle.logical_xor}[op]


def _bitwise(op):
    return {"and": paddle.bitwise_and, "or": paddle.bitwise_or, "xor": paddle.bitwise_xor}[op]


def test_compare_logical_bitwise_golden(gold):
    for name, _, _ in BCAST:
        for dt in INT_TYPES + ["float32", "float64", "bfloat16", "float16"]:
            x = to_dev(gold[f"fam_{name}_{dt}__in0"], HALF.get(dt))
            y = to_dev(gold[f"fam_{name}_{dt}__in1"], HALF.get(dt))
            for op in COMPARE_OPS:
                got = getattr(paddle, op)(x, y)
                assert got.dtype == torch.bool
                assert np.array_equal(to_np(got), gold[f"cmp_{op}_{name}_{dt}__out"]), (op, name, dt)
            if dt in HALF:
                continue
            if dt != "uint8":
                for op in ("and", "or", "xor"):
                    assert np.array_equal(to_np(_logical(op)(x, y)), gold[f"log_{op}_{name}_{dt}__out"]), (op, name, dt)
            if dt in INT_TYPES:
                for op in ("and", "or", "xor"):
                    got = _bitwise(op)(x, y)
                    assert got.dtype == x.dtype
                    assert np.array_equal(to_np(got), gold[f"bit_{op}_{name}_{dt}__out"]), (op, name, dt)


def test_unary_scale_cast_where_golden(gold):
    for dt in INT_TYPES + ["float32", "float64"]:
        x = to_dev(gold[f"un_{dt}__in0"])
        if dt != "uint8":
            assert np.array_equal(to_np(paddle.logical_not(x)), gold[f"lnot_{dt}__out"])
        if dt in INT_TYPES:
            assert np.array_equal(to_np(paddle.bitwise_not(x)), gold[f"bnot_{dt}__out"])
    for dt in ["uint8", "int8", "int16", "int32", "int64", "float32", "float64", "bool", "bfloat16"]:
        x = to_dev(gold[f"scale_{dt}__in0"], HALF.get(dt))
        for k, (s, b, after) in enumerate(SCALE_ARGS):
            got, want = to_np(paddle.scale(x, s, b, after)), gold[f"scale_{dt}_{k}__out"]
            if dt in ("float32", "float64"):
                np.testing.assert_allclose(got, want, rtol=2 * np.finfo(dt).eps, atol=np.finfo(dt).eps)
            elif dt == "bfloat16":
                xf = half_bits_to_f32(gold[f"scale_{dt}__in0"], "bf16")
                term = max(np.abs(s * xf).max(), abs(b), abs(s * b)) if xf.size else 0.0
                assert_half_close(got, want, 1, f"scale bf16 #{k}", "bf16", atol=float(term) * 2.0**-8)
            else:
                assert np.array_equal(got, want), (dt, k)
    for sc in range(10):
        x = to_dev(gold[f"cast_{sc}__in0"], CODE_HALF.get(sc))
        for dc in range(10):
            if dc == O.FLOAT16 and sc != O.FLOAT16:
                continue
            got = paddle.cast(x, CODE2TORCH[dc])
            assert got.dtype == CODE2TORCH[dc]
            assert np.array_equal(to_np(got), gold[f"cast_{sc}_{dc}__out"]), (sc, dc)
    for dt in ("int32", "int64", "float32", "float64"):
        c, x, y = (to_dev(gold[f"where_{dt}__in{i}"]) for i in range(3))
        assert np.array_equal(to_np(paddle.where(c, x, y)), gold[f"where_{dt}__out"])


# every launch route of the binary engine (same / inner / outer / general, aligned and not, with tails)
ROUTES = [((1 << 20,), (1 << 20,)), ((1000003,), (1000003,)), ((257, 4096), (4096,)), ((4096,), (257, 4096)),
          ((300, 1031), (1031,)), ((513, 2048), (513, 1)), ((513, 1), (513, 2048)), ((7, 1, 33, 16), (5, 1, 16)),
          ((65, 70000), (65, 1)), ((9, 10), ()), ((3, 1, 5), (1, 4, 1))]


@pytest.mark.parametrize("xs,ys", ROUTES)
def test_binary_routes(port, xs, ys):
    rng = np.random.default_rng(5)
    for dt in ("float32", "int64", "uint8", "bool", "int16", "float64"):
        x, y = draw(rng, dt, xs), draw(rng, dt, ys)
        tx, ty = to_dev(x), to_dev(y)
        for op in COMPARE_OPS:
            assert np.array_equal(to_np(getattr(paddle, op)(tx, ty)), port.compare(op, x, y)), (op, dt)
        if dt != "uint8":
            for op in ("and", "or", "xor"):
                assert np.array_equal(to_np(_logical(op)(tx, ty)), port.logical(op, x, y)), (op, dt)
        if dt in INT_TYPES:
            for op in ("and", "or", "xor"):
                assert np.array_equal(to_np(_bitwise(op)(tx, ty)), port.bitwise(op, x, y)), (op, dt)
    for half, conv, code in (("bf16", port.f32_to_bf16, O.BFLOAT16), ("fp16", port.f32_to_fp16, O.FLOAT16)):
        x = conv(draw(rng, "float32", xs).astype(np.float32))
        y = conv(draw(rng, "float32", ys).astype(np.float32))
        tx, ty = to_dev(x, half), to_dev(y, half)
        for op in COMPARE_OPS:
            assert np.array_equal(to_np(getattr(paddle, op)(tx, ty)), port.compare(op, x, y, code=code)), (op, half)
        for op in ("and", "or", "xor"):  # GPU-only registration: the same functor on the 16-bit value
            assert np.array_equal(to_np(_logical(op)(tx, ty)), port.logical(op, x, y, code=code)), (op, half)
        assert np.array_equal(to_np(paddle.logical_not(tx)), port.logical_not(x, code=code))


def test_unaligned_views_and_tails(port):
    """Operands that start in the middle of an allocation (not 16-byte aligned) and odd lengths."""
    rng = np.random.default_rng(6)
    for dt, tdt in (("float32", torch.float32), ("int16", torch.int16), ("uint8", torch.uint8)):
        n = 100003
        x, y = draw(rng, dt, (n + 3,)), draw(rng, dt, (n + 3,))
        tx, ty = to_dev(x)[1:n + 1], to_dev(y)[3:n + 3]
        assert np.array_equal(to_np(paddle.less_than(tx, ty)), port.compare("less_than", x[1:n + 1], y[3:n + 3]))
        assert np.array_equal(to_np(paddle.equal(tx, ty)), port.compare("equal", x[1:n + 1], y[3:n + 3]))
        got = paddle.cast(tx, torch.float64)
        assert np.array_equal(to_np(got), port.cast(x[1:n + 1], O.FLOAT64), equal_nan=dt == "float32")
        if dt != "float32":
            assert np.array_equal(to_np(paddle.bitwise_not(tx)), port.bitwise_not(np.ascontiguousarray(x[1:n + 1])))
        c = rng.random(n + 3) < 0.5
        tc = to_dev(c)[2:n + 2]
        w_got, w_want = to_np(paddle.where(tc, tx, ty)), port.where(c[2:n + 2], x[1:n + 1], y[3:n + 3])
        assert np.array_equal(w_got.view(np.uint8), w_want.view(np.uint8))  # a bit select: NaNs included


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 4099, 1 << 21])
def test_unary_sizes(port, n):
    rng = np.random.default_rng(n + 1)
    xi = rng.integers(-2**40, 2**40, n)
    assert np.array_equal(to_np(paddle.scale(to_dev(xi), 3.0, -7.0, False)), port.scale(xi, 3.0, -7.0, False))
    assert np.array_equal(to_np(paddle.bitwise_not(to_dev(xi))), port.bitwise_not(xi))
    assert np.array_equal(to_np(paddle.cast(to_dev(xi), "int16")), port.cast(xi, O.INT16))
    xf = (rng.standard_normal(n) * 40).astype(np.float32)
    np.testing.assert_allclose(to_np(paddle.scale(to_dev(xf), 0.1, -0.3)), port.scale(xf, 0.1, -0.3), rtol=3e-7, atol=2e-7)
    assert np.array_equal(to_np(paddle.cast(to_dev(xf), "bfloat16")), port.cast(xf, O.BFLOAT16))
    assert np.array_equal(to_np(paddle.cast(to_dev(xf), "float16")), port.cast(xf, O.FLOAT16))  # both round to nearest even
    assert np.array_equal(to_np(paddle.cast(to_dev(xf), "int32")), port.cast(xf, O.INT32))
    assert np.array_equal(to_np(paddle.logical_not(to_dev(xf))), port.logical_not(xf))
    for half, conv, code in (("bf16", port.f32_to_bf16, O.BFLOAT16), ("fp16", port.f32_to_fp16, O.FLOAT16)):
        xb = conv(xf)
        for s, b, after in ((2.0, 1.0, True), (0.1, -0.3, True), (3.0, 0.5, False)):
            got, want = to_np(paddle.scale(to_dev(xb, half), s, b, after)), port.scale(xb, s, b, after, code=code)
            term = float(max(np.abs(s * xf).max(), abs(b), abs(s * b))) if n else 0.0
            assert_half_close(got, want, 1, f"scale {half}", half, atol=term * (2.0**-8 if half == "bf16" else 2.0**-11))
        for dc in (O.FLOAT32, O.FLOAT64, O.INT64, O.BOOL, O.UINT8 if n == 0 else O.INT8):
            assert np.array_equal(to_np(paddle.cast(to_dev(xb, half), CODE2TORCH[dc])), port.cast(xb, dc, code=code))


def test_where_all_sizes(port):
    rng = np.random.default_rng(8)
    for tdt, dt in ((torch.bool, np.bool_), (torch.int8, np.int8), (torch.int16, np.int16), (torch.float32, np.float32),
                    (torch.int64, np.int64), (torch.float64, np.float64)):
        for shape in ((5, 7), (1 << 20,), (333, 1001)):
            c = rng.random(shape) < 0.5
            x = (rng.standard_normal(shape) * 50).astype(dt)
            y = (rng.standard_normal(shape) * 50).astype(dt)
            assert np.array_equal(to_np(paddle.where(to_dev(c), to_dev(x), to_dev(y))), port.where(c, x, y))
    xb, yb = port.f32_to_bf16(rng.standard_normal(4099)), port.f32_to_bf16(rng.standard_normal(4099))
    c = rng.random(4099) < 0.5
    assert np.array_equal(to_np(paddle.where(to_dev(c), to_dev(xb, "bf16"), to_dev(yb, "bf16"))), np.where(c, xb, yb))


def test_bitwise_shift_golden_and_routes(port, gold):
    from golden.make_golden_family import SHIFT_CASES
    fns = {0: paddle.bitwise_left_shift, 1: paddle.bitwise_right_shift}
    for name, _, _ in SHIFT_CASES:
        for dt in ("uint8", "int8", "int16", "int32", "int64"):
            x, y = to_dev(gold[f"sh_{name}_{dt}__in0"]), to_dev(gold[f"sh_{name}_{dt}__in1"])
            for right in (0, 1):
                for ar in (0, 1):
                    assert np.array_equal(to_np(fns[right](x, y, is_arithmetic=bool(ar))), gold[f"sh_{name}_{dt}_{right}{ar}__out"])
    rng = np.random.default_rng(9)
    for xs, ys in ROUTES[:8]:
        for dt in (np.int32, np.int64, np.uint8):
            info = np.iinfo(dt)
            x = rng.integers(info.min, info.max, xs, dtype=dt, endpoint=True)
            y = rng.integers(0 if dt == np.uint8 else -2, info.bits + 2, ys).astype(dt)
            for right in (0, 1):
                for ar in (True, False):
                    assert np.array_equal(to_np(fns[right](to_dev(x), to_dev(y), is_arithmetic=ar)), port.bitwise_shift(right, x, y, ar))


def test_family_errors():
    x = torch.zeros(4, 5, device="cuda")
    with pytest.raises(ValueError, match="Broadcast dimension mismatch|broadcast"):
        paddle.less_than(x, torch.zeros(4, 6, device="cuda"))
    with pytest.raises(NotImplementedError, match="not registered"):
        paddle.logical_and(x.to(torch.uint8), x.to(torch.uint8))
    with pytest.raises(NotImplementedError, match="not registered"):
        paddle.bitwise_and(x, x)
    with pytest.raises(NotImplementedError, match="not registered"):
        paddle.bitwise_not(x)
    with pytest.raises(ValueError, match="same shape"):
        paddle.where(x > 0, x, torch.zeros(5, device="cuda"))
    with pytest.raises(TypeError):
        paddle.where(x, x, x)
    with pytest.raises(RuntimeError, match="no CPU kernels"):
        paddle.scale(torch.zeros(3), 2.0)


def test_config2_sized_properties():
    """BASELINE config 2 shapes (x[64,2048,4096] bf16 against y[4096]): exactly one of x < y and x >= y holds for every
    element without a NaN, checked with this library's own reductions; where() picks from the right side."""
    g = torch.Generator(device="cuda")
    g.manual_seed(2)
    x = torch.randn(64, 2048, 4096, device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randn(4096, device="cuda", generator=g).to(torch.bfloat16)
    lt, ge = paddle.less_than(x, y), paddle.greater_equal(x, y)
    assert int(paddle.sum(paddle.logical_xor(lt, ge)).item()) == x.numel()
    assert int(paddle.sum(paddle.logical_and(lt, ge)).item()) == 0
    n_lt = int(paddle.sum(lt).item())
    assert 0.3 * x.numel() < n_lt < 0.7 * x.numel()
    yb = paddle.add(paddle.scale(x, 0.0), y)  # y broadcast to x's shape with this library's own kernels
    m = paddle.where(lt, x, yb)               # = minimum(x, y)
    assert int(paddle.sum(paddle.not_equal(m, paddle.minimum(x, y))).item()) == 0
    back = paddle.cast(paddle.cast(x, "float32"), "bfloat16")
    assert int(paddle.sum(paddle.not_equal(back, x)).item()) == 0


def test_bf16_divide_is_exact_for_every_pair_of_operands():
    """The bf16 divide skips the IEEE division sequence (approximate quotient, then one rounding — see DivOp<bf16>).
    Checked here against round_bf16(float(a) / float(b)) with the exact fp32 division, for ALL 2^32 operand pairs,
    bit for bit (NaN payloads included), using this library's own fp32 divide, cast and integer compare."""
    allbits = torch.arange(65536, device="cuda", dtype=torch.int32).to(torch.int16)  # every bf16 bit pattern
    y = allbits.view(torch.bfloat16)
    yf = paddle.cast(y, "float32")
    bad = 0
    for lo in range(0, 65536, 8192):
        x = allbits[lo:lo + 8192].view(torch.bfloat16).reshape(8192, 1)
        fast = paddle.divide(x, y)
        exact = paddle.cast(paddle.divide(paddle.cast(x, "float32"), yf), "bfloat16")
        bad += int(paddle.sum(paddle.not_equal(fast.view(torch.int16), exact.view(torch.int16))).item())
        del fast, exact
    assert bad == 0, f"{bad} of 2^32 bf16 quotients differ from the correctly rounded result"


@pytest.mark.parametrize("half,code", [("bf16", O.BFLOAT16), ("fp16", O.FLOAT16)])
def test_half_equal_not_equal_every_value_against_special_operands(port, half, code):
    """The packed 16-bit equal / not_equal (csrc/elementwise_pred.cu: eq8_packed) against the scalar rule of the CPU functor
    (|round16(a - b)| < 1e-8, infinities exact, NaN equals nothing): EVERY 16-bit pattern as `a`, against operands chosen
    where the rule has corners — zeros of both signs, infinities, NaNs, subnormals, values around the 1e-8 tolerance,
    neighbours of `a` itself — on the same-shape (vector) route and, for one operand, the row-broadcast route."""
    a = np.arange(65536, dtype=np.uint16)
    conv = port.f32_to_bf16 if half == "bf16" else port.f32_to_fp16
    special = np.array([0x0000, 0x8000, 0x7C00, 0xFC00, 0x7F80, 0xFF80, 0x7FC0, 0xFFC0, 0x7E00, 0x0001, 0x8001, 0x0002, 0x007F, 0x0080,
                        0x03FF, 0x0400, 0x3C00, 0x3F80, 0xBF80, 0x7BFF, 0x7F7F], dtype=np.uint16)
    tiny = conv(np.array([1e-8, 9.9e-9, 1.01e-8, 5e-9, 2e-9, 1e-9, -1e-9, -5e-9, 1.5e-8, 2e-8, 6e-8, 1e-7, 1e-10, 3e-39],
                         dtype=np.float32))
    bs = np.concatenate([special, tiny])
    A = np.tile(a, (len(bs) + 3, 1))
    B = np.empty_like(A)
    for i, b in enumerate(bs):
        B[i] = b
    B[len(bs)] = a                      # a == a (NaN patterns must still be unequal)
    B[len(bs) + 1] = np.roll(a, 1)      # the neighbouring bit pattern
    B[len(bs) + 2] = a ^ np.uint16(0x8000)  # the same magnitude, other sign
    ta, tb = to_dev(A, half), to_dev(B, half)
    for op, fn in (("equal", paddle.equal), ("not_equal", paddle.not_equal)):
        want = port.compare(op, A, B, code=code)
        got = to_np(fn(ta, tb))
        bad = np.argwhere(got != want)
        assert bad.size == 0, (op, half, [(hex(A[i, j]), hex(B[i, j]), bool(got[i, j])) for i, j in bad[:5]])
    # row-broadcast route: every `a` row against one row of operands
    row = np.resize(bs, 65536).astype(np.uint16)
    want = port.compare("equal", A[:4], row, code=code)
    assert np.array_equal(to_np(paddle.equal(to_dev(A[:4], half), to_dev(row, half))), want)
