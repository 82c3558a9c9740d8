"""GPU parity for max_grad / min_grad / prod_grad (the remaining reduction backward kernels) through
the C-ABI.  Bit-exact against the compiled reference CPU kernels for the types they register (float32 / float64 / int32 /
int64) on inputs where the GPU kernel's EqualFunctor (|a - b| < 1e-8) and the CPU kernel's exact == agree, and against the
restatement with the GPU rule for bf16 / fp16 and for values that differ by less than the tolerance."""
import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

ROUTES = [((4096, 512), 0), ((4096, 512), 1), ((4096, 512), None), ((1000, 264), 0), ((257, 1031), 0), ((257, 1031), 1),
          ((64, 8, 128), 1), ((64, 8, 128), (0, 2)), ((7, 33, 5, 16), (1, 3)), ((1 << 18,), 0), ((3, 70000), 1), ((65, 1, 9), 1),
          ((2, 3, 4, 5, 6), (0, 2, 4)), ((5, 4096), 0), ((300, 40), 1), ((), None),
          ((16, 13, 2048), 1), ((3, 70, 1024), 1)]  # one middle dimension, long inner rows: the batched inner route
FNS = {"max": paddle.max_grad, "min": paddle.min_grad, "prod": paddle.prod_grad}


def _axes(axis):
    return None if axis is None else ([axis] if isinstance(axis, int) else list(axis))


@pytest.mark.parametrize("shape,axis", ROUTES)
@pytest.mark.parametrize("keep", [False, True])
def test_select_grad_routes_against_compiled_reference(ref, shape, axis, keep):
    rng = np.random.default_rng(11)
    axes = _axes(axis)
    for dt in (np.float32, np.int64, np.float64, np.int32):
        flt = np.issubdtype(dt, np.floating)
        x = rng.integers(-4, 5, shape).astype(dt)  # ties in almost every slice
        for op, fwd in (("max", np.max), ("min", np.min)):
            out = np.asarray(fwd(x, axis=axis if axis is None else tuple(axes), keepdims=keep)).astype(dt)
            g = (rng.standard_normal(out.shape) * 100).astype(dt)
            got = FNS[op](to_dev(x).reshape(shape), to_dev(out).reshape(out.shape), to_dev(g).reshape(out.shape), axis=axis, keepdim=keep)
            assert np.array_equal(to_np(got).reshape(-1), ref.select_grad(op, x, out, g, axes, keep).reshape(-1)), (op, dt)
        xp = x.copy()
        xp[xp == 0] = 3
        if flt:
            xp = (xp * rng.uniform(0.5, 1.5, shape)).astype(dt)
        with np.errstate(over="ignore"):
            out = np.asarray(np.prod(xp, axis=axis if axis is None else tuple(axes), keepdims=keep)).astype(dt)
        if flt:
            out = np.clip(np.nan_to_num(out, nan=1.0), -1e30, 1e30).astype(dt)  # any "out" is a valid input of the kernel
        g = rng.standard_normal(out.shape).astype(dt) if flt else rng.integers(-5, 6, out.shape).astype(dt)
        got = paddle.prod_grad(to_dev(xp).reshape(shape), to_dev(out).reshape(out.shape), to_dev(g).reshape(out.shape), axis=axis, keepdim=keep)
        assert np.array_equal(to_np(got).reshape(-1), ref.select_grad("prod", xp, out, g, axes, keep).reshape(-1)), ("prod", dt)


@pytest.mark.parametrize("shape,axis", [((4096, 512), 0), ((4096, 512), 1), ((64, 8, 128), 1), ((257, 1031), 1), ((1 << 16,), None),
                                        ((16, 13, 2048), 1)])
def test_select_grad_16bit_against_restatement(port, shape, axis):
    rng = np.random.default_rng(12)
    axes = _axes(axis)
    for half, conv, back, code in (("bf16", port.f32_to_bf16, port.bf16_to_f32, O.BFLOAT16), ("fp16", port.f32_to_fp16, port.fp16_to_f32, O.FLOAT16)):
        xf = rng.integers(-4, 5, shape).astype(np.float32) * 0.25
        x = conv(xf)
        oshape = tuple(O.reduced_shape(shape, axes or [], False))
        for op, fwd in (("max", np.max), ("min", np.min)):
            out = conv(np.asarray(fwd(back(x), axis=axis if axis is None else tuple(axes))).astype(np.float32)).reshape(oshape)
            g = conv(rng.standard_normal(oshape).astype(np.float32)).reshape(oshape)
            got = FNS[op](to_dev(x, half).reshape(shape), to_dev(out, half).reshape(out.shape), to_dev(g, half).reshape(out.shape), axis=axis)
            assert np.array_equal(to_np(got).reshape(shape), port.select_grad(op, x, out, g, axes, code=code)), (op, half)
        xp = conv((rng.uniform(0.5, 2.0, shape) * rng.choice([-1.0, 1.0], shape)).astype(np.float32))
        out = conv(rng.standard_normal(oshape).astype(np.float32)).reshape(oshape)
        g = conv(rng.standard_normal(oshape).astype(np.float32)).reshape(oshape)
        got = paddle.prod_grad(to_dev(xp, half).reshape(shape), to_dev(out, half).reshape(out.shape), to_dev(g, half).reshape(out.shape), axis=axis)
        assert np.array_equal(to_np(got).reshape(shape), port.select_grad("prod", xp, out, g, axes, code=code)), ("prod", half)


def test_select_grad_tolerance_and_special_values(port):
    """The GPU rule: values closer than 1e-8 count as equal (the CPU kernel would not), infinities compare exactly, NaN
    equals nothing, and out_grad = inf makes the non-selected positions NaN (inf * 0)."""
    x = np.array([[1e-9, 2e-9, 5e-8, -3e-9], [np.inf, np.nan, np.inf, 1.0], [np.nan, np.nan, -np.inf, 0.0]], np.float32)
    out = np.array([5e-8, np.inf, 0.0], np.float32)
    for g in (np.array([2.0, 3.0, 4.0], np.float32), np.array([np.inf, 3.0, np.nan], np.float32)):
        got = to_np(paddle.max_grad(to_dev(x), to_dev(out), to_dev(g), axis=1))
        assert np.array_equal(got, port.select_grad("max", x, out, g, [1], tol=True), equal_nan=True)
    x2 = np.array([[1e-9, 2e-9, 9e-9, 1.2e-8]], np.float32)
    out2 = np.array([1e-9], np.float32)
    g2 = np.array([7.0], np.float32)
    got = to_np(paddle.min_grad(to_dev(x2), to_dev(out2), to_dev(g2), axis=1))
    assert np.array_equal(got, np.array([[7.0, 7.0, 7.0, 0.0]], np.float32))  # 8e-9 < 1e-8 < 1.1e-8
    assert np.array_equal(got, port.select_grad("min", x2, out2, g2, [1], tol=True))
    xd = x2.astype(np.float64)
    got = to_np(paddle.min_grad(to_dev(xd), to_dev(out2.astype(np.float64)), to_dev(g2.astype(np.float64)), axis=1))
    assert np.array_equal(got, port.select_grad("min", xd, out2.astype(np.float64), g2.astype(np.float64), [1], tol=True))
    # prod_grad with zeros and infinities in x: 1/0 = inf, (dy*y) * inf, 0 * inf = NaN — whatever IEEE says, step by step
    xz = np.array([[0.0, -0.0, np.inf, 2.0, np.nan, 1e-40]], np.float32)
    oz = np.array([0.0], np.float32)
    for gz in (np.array([1.0], np.float32), np.array([np.inf], np.float32)):
        got = to_np(paddle.prod_grad(to_dev(xz), to_dev(oz), to_dev(gz), axis=1))
        assert np.array_equal(got, port.select_grad("prod", xz, oz, gz, [1]), equal_nan=True)
    xi = np.array([[0, 1, -1, 2, -7, 1]], np.int32)  # integer 1/x: 0 (defined here for x == 0), 1, -1, 0, 0, 1
    got = to_np(paddle.prod_grad(to_dev(xi), to_dev(np.array([0], np.int32)), to_dev(np.array([5], np.int32)), axis=1))
    assert np.array_equal(got, np.zeros((1, 6), np.int32))
    got = to_np(paddle.prod_grad(to_dev(xi), to_dev(np.array([3], np.int32)), to_dev(np.array([5], np.int32)), axis=1))
    assert np.array_equal(got, np.array([[0, 15, -15, 0, 0, 15]], np.int32))


def test_prod_grad_reciprocal_every_16bit_pattern(port):
    """The 16-bit 1 / x takes the approximate reciprocal where its error cannot change the rounding (csrc/reduce_grad.cu):
    with out = out_grad = 1 the result is round(1 / x); every bit pattern of both types, against the exact division."""
    bits = np.arange(65536, dtype=np.uint32).astype(np.uint16).reshape(256, 256)
    one = {"bf16": 0x3F80, "fp16": 0x3C00}
    for half, code in (("bf16", O.BFLOAT16), ("fp16", O.FLOAT16)):
        ones = np.full((256,), one[half], np.uint16)
        got = to_np(paddle.prod_grad(to_dev(bits, half), to_dev(ones, half), to_dev(ones, half), axis=1))
        want = port.select_grad("prod", bits, ones, ones, [1], code=code)
        nan = port.bf16_to_f32(want) != port.bf16_to_f32(want) if half == "bf16" else port.fp16_to_f32(want) != port.fp16_to_f32(want)
        gf = port.bf16_to_f32(got) if half == "bf16" else port.fp16_to_f32(got)
        assert np.array_equal(got[~nan], want[~nan]), half
        assert np.all(gf[nan] != gf[nan]), half  # NaN stays NaN (the payload is not compared)


def test_select_grad_unaligned_and_errors():
    x = torch.randn(77 * 1031 + 1, device="cuda")[1:].reshape(77, 1031)  # not 16-byte aligned
    out = x.max(dim=0).values
    g = torch.randn(1031, device="cuda")
    got = paddle.max_grad(x, out, g, axis=0)
    assert torch.equal(got, torch.where(x == out, g.expand(77, 1031), torch.zeros((), device="cuda")))
    with pytest.raises(ValueError, match="has shape"):
        paddle.max_grad(x, out[:5], g, axis=0)
    with pytest.raises(ValueError):
        paddle.prod_grad(x, out, g, axis=2)
    with pytest.raises(TypeError):
        paddle.min_grad(x, out.double(), g, axis=0)
    xi = torch.zeros((4, 5), device="cuda", dtype=torch.int8)
    with pytest.raises(NotImplementedError, match="not registered"):
        paddle.max_grad(xi, xi[0], xi[0], axis=0)


def test_round_trip_with_forward_at_config_sizes():
    """Size-independent properties at a config-3-sized tensor, using this library's own forward kernels: with distinct
    values per row, max_grad has exactly one non-zero per row and its row sums give out_grad back; max_grad + (x < max)
    partition every row; prod_grad * x == out_grad * out wherever the arithmetic is exact (powers of two)."""
    rows, cols = 32768, 8192
    x = torch.rand(rows, cols, device="cuda")
    col = torch.randint(0, cols, (rows,), device="cuda")
    x[torch.arange(rows, device="cuda"), col] = 2.0  # the unique maximum of each row
    out = paddle.max(x, axis=-1)
    g = torch.randn(rows, device="cuda")
    dx = paddle.max_grad(x, out, g, axis=-1)
    assert torch.equal(paddle.sum(dx, axis=-1), g)  # one non-zero term per row: exact
    assert torch.equal(paddle.argmax(paddle.not_equal(dx, torch.zeros((), device="cuda")).to(torch.int32), axis=-1), col)
    del dx
    e = torch.randint(-3, 4, (rows, cols), device="cuda").float()
    xp = torch.exp2(e)
    outp = torch.exp2(torch.randint(-8, 9, (rows,), device="cuda").float())
    dp = paddle.prod_grad(xp, outp, g, axis=-1)
    assert torch.equal(paddle.multiply(dp, xp), (g * outp)[:, None].expand(rows, cols))
