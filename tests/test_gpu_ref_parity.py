"""GPU kernels against the COMPILED REFERENCE (oracle/_ref/libphi_cpu_ref.so: Paddle's own CPU PHI kernels, built by
oracle/build_ref.sh and shipped to the GPU box prebuilt) — no restatement in between.  One test per op of the hot path,
at sizes that take the routes the BASELINE configs take (TMA row pipeline, interleaved split reduce, TMA column
tiles), plus the special-value rows the routes must agree on.  Skipped only if the reference library was not shipped.
Bars: bit-exact for integer / max / min / argmax / add / multiply / compare / cast / where; rtol 1e-5 for fp32
reductions, softmax, layer_norm; 1 ulp for 16-bit outputs."""
import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional


def _rng(seed):
    return np.random.default_rng(seed)


def test_ref_add_multiply_bf16_int32(ref):
    r = _rng(1)
    a = ref.f32_to_bf16(r.standard_normal((512, 4096)).astype(np.float32))
    b = ref.f32_to_bf16(r.standard_normal((4096,)).astype(np.float32))
    for op, fn in (("add", paddle.add), ("multiply", paddle.multiply)):
        # (the CPU add kernel is not registered for bf16: the shim runs ElementwiseCompute with bfloat16's own operator+)
        want = ref.elementwise(op, a, b, O.BFLOAT16)
        assert np.array_equal(to_np(fn(to_dev(a, "bf16"), to_dev(b, "bf16"))), want), op
    ai = r.integers(-2**31, 2**31, (512, 4096), dtype=np.int32)
    bi = r.integers(-2**31, 2**31, (4096,), dtype=np.int32)
    for op, fn in (("add", paddle.add), ("multiply", paddle.multiply), ("subtract", paddle.subtract)):
        assert np.array_equal(to_np(fn(to_dev(ai), to_dev(bi))), ref.elementwise(op, ai, bi)), op
    # general N-D broadcast route
    x5 = r.integers(-100, 100, (6, 1, 8, 1, 32), dtype=np.int64)
    y5 = r.integers(-100, 100, (1, 5, 1, 7, 32), dtype=np.int64)
    assert np.array_equal(to_np(paddle.add(to_dev(x5), to_dev(y5))), ref.elementwise("add", x5, y5))


@pytest.mark.parametrize("shape,axis", [((4096, 4096), -1), ((300, 8192), 0), ((64, 96, 128), 1), ((1 << 22,), None),
                                         ((33, 1001), [0, 1])])
def test_ref_reductions(ref, shape, axis):
    r = _rng(2)
    xf = r.random(shape, dtype=np.float32) * 2 - 1
    xi = r.integers(-2**40, 2**40, shape, dtype=np.int64)
    tf, ti = to_dev(xf), to_dev(xi)
    absum = np.abs(xf).sum(axis=tuple(axis) if isinstance(axis, list) else axis)
    for op, fn in (("sum", paddle.sum), ("mean", paddle.mean)):
        got, want = to_np(fn(tf, axis=axis)), ref.reduce(op, xf, axis)
        scale = absum if op == "sum" else absum / (xf.size / max(1, want.size))
        assert np.all(np.abs(got - want) <= 1e-5 * np.maximum(scale, 1e-30)), (op, shape, axis)
    for op, fn in (("max", paddle.max), ("min", paddle.min)):
        assert np.array_equal(to_np(fn(tf, axis=axis)), ref.reduce(op, xf, axis)), (op, shape, axis)
        assert np.array_equal(to_np(fn(ti, axis=axis)), ref.reduce(op, xi, axis)), (op, shape, axis)
    assert np.array_equal(to_np(paddle.sum(ti, axis=axis)), ref.reduce("sum", xi, axis)), ("sum int64", shape, axis)
    small = r.integers(-3, 4, shape, dtype=np.int64)
    assert np.array_equal(to_np(paddle.prod(to_dev(small), axis=axis)), ref.reduce("prod", small, axis)), ("prod", shape)


def test_ref_argmax(ref):
    r = _rng(3)
    x = r.integers(0, 50, (257, 4096), dtype=np.int64)  # many ties: lowest index must win
    for axis in (None, 0, 1, -1):
        assert np.array_equal(to_np(paddle.argmax(to_dev(x), axis=axis)), ref.argmax(x, axis)), axis
    xf = r.standard_normal((64, 70000)).astype(np.float32)
    assert np.array_equal(to_np(paddle.argmax(to_dev(xf), axis=1)), ref.argmax(xf, 1))
    # 1- and 2-byte values with inner % 16 == 0 on the outer axes: the vector column kernel (16 columns x 16-byte
    # (value, index) pairs per thread = 64 KB of shared memory for uint8: needs the opt-in)
    for dt in (np.uint8, np.int16):
        info = np.iinfo(dt)
        xs = r.integers(info.min, info.max, (6, 40, 64), dtype=dt, endpoint=True)
        for axis in (0, 1, 2):
            assert np.array_equal(to_np(paddle.argmax(to_dev(xs), axis=axis)), ref.argmax(xs, axis)), (dt, axis)
        big = r.integers(info.min, info.max, (3, 300, 4096), dtype=dt, endpoint=True)
        for axis in (0, 1):
            assert np.array_equal(to_np(paddle.argmax(to_dev(big), axis=axis)), ref.argmax(big, axis)), (dt, axis, "big")


def test_ref_softmax_pipe_and_cluster_sizes(ref):
    r = _rng(4)
    for shape in ((4096, 2048), (1024, 8192), (8, 128256), (600, 1000)):
        x = (r.standard_normal(shape) * 2).astype(np.float32)
        np.testing.assert_allclose(to_np(F.softmax(to_dev(x), axis=-1)), ref.softmax(x, -1), rtol=1e-5, atol=1e-30)
        np.testing.assert_allclose(to_np(F.log_softmax(to_dev(x), axis=-1)), ref.log_softmax(x, -1), rtol=1e-5, atol=1e-5)
    x = (r.standard_normal((32, 48, 64)) * 2).astype(np.float32)
    np.testing.assert_allclose(to_np(F.softmax(to_dev(x), axis=1)), ref.softmax(x, 1), rtol=1e-5, atol=1e-30)


@pytest.mark.parametrize("shape", [(4096, 2048), (2048, 8192), (8, 128256), (64, 512)])
@pytest.mark.parametrize("dt", ["float32", "bfloat16"])
def test_softmax_special_values_every_route(port, shape, dt):
    """NaN / +inf / all -inf / huge-outlier rows at pipe-, cluster- and register-route sizes: every route must
    PROPAGATE — a row holding a NaN or a +inf, or made of -inf only, comes out all NaN — and clip a -1e4 outlier to
    exp(-64).  That is the rule of the reference's scalar clip `x < -64 ? -64 : x` (funcs/cpu_vec.h:383-387, ValueClip
    funcs/softmax_impl.h:36-42), of its float64 route and of its GPU kernels.  The reference's float32 CPU route is not
    usable as the oracle for these rows: its AVX clip `_mm256_max_ps(x, threshold)` (funcs/cpu_vec.h:406) returns the
    threshold for a NaN operand, so on an AVX host a NaN logit in the first n - n % 8 columns comes out as exp(-64) and
    the rest of the row looks healthy, while the same value in the scalar tail propagates.  The restatement follows the
    scalar rule everywhere; it is the expected value here."""
    r = _rng(5)
    x = (r.standard_normal(shape) * 2).astype(np.float32)
    x[1, 7] = np.nan
    x[2, shape[1] - 1] = np.inf
    x[3, :] = -np.inf
    x[4, 5] = -1e4
    x[5, 0] = -np.inf
    if dt == "bfloat16":
        xb = port.f32_to_bf16(x)
        want32 = port.softmax(port.bf16_to_f32(xb), -1)
        got = port.bf16_to_f32(to_np(F.softmax(to_dev(xb, "bf16"), axis=-1)))
        tol = dict(rtol=2 ** -7, atol=1e-30)
    else:
        want32 = port.softmax(x, -1)
        got = to_np(F.softmax(to_dev(x), axis=-1))
        tol = dict(rtol=1e-5, atol=2e-28)
    for row in (1, 2, 3):
        assert np.isnan(want32[row]).all(), "oracle sanity"
        assert np.isnan(got[row]).all(), f"row {row} must be NaN, got finite values (NaN dropped by the kernel)"
    keep = [i for i in range(min(shape[0], 64)) if i not in (1, 2, 3)]
    np.testing.assert_allclose(got[keep], want32[keep], **tol)
    lg = to_np(F.log_softmax(to_dev(x), axis=-1)) if dt == "float32" else None
    if lg is not None:
        for row in (1, 2, 3):
            assert np.isnan(lg[row]).all(), f"log_softmax row {row} must be NaN"
        np.testing.assert_allclose(lg[keep], port.log_softmax(x, -1)[keep], rtol=1e-5, atol=1e-5)


def test_ref_layer_norm_pipe_and_reg(ref):
    r = _rng(6)
    for rows, D in ((4096, 8192), (2048, 1024), (700, 768), (16, 40000)):
        x = (r.standard_normal((rows, D)) * 3 + 1).astype(np.float32)
        w = r.uniform(0.5, 1.5, D).astype(np.float32)
        b = r.uniform(-0.5, 0.5, D).astype(np.float32)
        oy, om, ov = ref.layer_norm(x, 1, w, b, 1e-5)
        y, m, v = F.layer_norm(to_dev(x), [D], to_dev(w), to_dev(b), 1e-5, return_stats=True)
        np.testing.assert_allclose(to_np(y), oy, rtol=1e-5, atol=2e-5)
        np.testing.assert_allclose(to_np(m), om, rtol=1e-5, atol=1e-5 * float(np.abs(x).max()))
        np.testing.assert_allclose(to_np(v), ov, rtol=2e-5)
        xb = ref.f32_to_bf16(x)
        oy, om, ov = ref.layer_norm(ref.bf16_to_f32(xb), 1, w, b, 1e-5)
        yb = F.layer_norm(to_dev(xb, "bf16"), [D], to_dev(w), to_dev(b), 1e-5)
        assert_half_close(to_np(yb), ref.f32_to_bf16(oy), 1, f"layer_norm bf16 [{rows},{D}]", atol=2e-5)


def test_ref_backward_ops(ref):
    r = _rng(7)
    y = ref.softmax((r.standard_normal((4096, 2048)) * 2).astype(np.float32), -1)
    dy = r.standard_normal((4096, 2048)).astype(np.float32)
    np.testing.assert_allclose(to_np(F.softmax_grad(to_dev(y), to_dev(dy), axis=-1)), ref.softmax_grad(y, dy, -1), rtol=1e-5,
                               atol=1e-6)
    rows, D = 2048, 4096
    x = (r.standard_normal((rows, D)) * 2 + 0.5).astype(np.float32)
    w = r.uniform(0.5, 1.5, D).astype(np.float32)
    g = r.standard_normal((rows, D)).astype(np.float32)
    _, m, v = ref.layer_norm(x, 1, w, None, 1e-5)
    dx, ds, db = ref.layer_norm_grad(x, 1, w, m, v, g, 1e-5)
    gx, gs, gb = F.layer_norm_grad(to_dev(x), [D], to_dev(w), to_dev(m), to_dev(v), to_dev(g), 1e-5)
    np.testing.assert_allclose(to_np(gx), dx, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(to_np(gs), ds, rtol=1e-4, atol=1e-4 * float(np.abs(ds).max()))
    np.testing.assert_allclose(to_np(gb), db, rtol=1e-4, atol=1e-4 * float(np.abs(db).max()))
    a = r.integers(-2**31, 2**31, (600, 1024), dtype=np.int64)
    b = r.integers(-2**31, 2**31, (1024,), dtype=np.int64)
    d = r.integers(-2**31, 2**31, (600, 1024), dtype=np.int64)
    for op, fn in (("add", paddle.add_grad), ("multiply", paddle.multiply_grad)):
        wx, wy = ref.elementwise_grad(op, a, b, d)
        qx, qy = fn(to_dev(a), to_dev(b), to_dev(d))
        assert np.array_equal(to_np(qx), wx) and np.array_equal(to_np(qy), wy), op
    og = r.standard_normal((64,)).astype(np.float32)
    assert np.array_equal(to_np(paddle.sum_grad(to_dev(np.zeros((64, 96), np.float32)), to_dev(og), axis=1)),
                          ref.reduce_grad("sum", (64, 96), og, 1))


def test_ref_family(ref):
    r = _rng(8)
    a = ref.f32_to_bf16(r.standard_normal((300, 4096)).astype(np.float32))
    b = ref.f32_to_bf16(r.standard_normal((4096,)).astype(np.float32))
    ta, tb = to_dev(a, "bf16"), to_dev(b, "bf16")
    for op, fn in (("less_than", paddle.less_than), ("greater_equal", paddle.greater_equal), ("equal", paddle.equal),
                   ("not_equal", paddle.not_equal)):
        assert np.array_equal(to_np(fn(ta, tb)), ref.compare(op, a, b, code=O.BFLOAT16)), op
    for op, fn in (("divide", paddle.divide), ("maximum", paddle.maximum)):
        assert np.array_equal(to_np(fn(ta, tb)), ref.elementwise(op, a, b, O.BFLOAT16)), op
    x = (r.standard_normal((1 << 20,)) * 100).astype(np.float32)
    assert np.array_equal(to_np(paddle.cast(to_dev(x), "bfloat16")), ref.cast(x, O.BFLOAT16))
    assert np.array_equal(to_np(paddle.cast(to_dev(x), "int32")), ref.cast(x, O.INT32))
    c = r.random(1 << 20) < 0.5
    y = r.standard_normal(1 << 20).astype(np.float32)
    assert np.array_equal(to_np(paddle.where(to_dev(c), to_dev(x), to_dev(y))), ref.where(c, x, y))
    xi = r.integers(-2**31, 2**31, (1 << 20,), dtype=np.int32)
    yi = r.integers(-2**31, 2**31, (1 << 20,), dtype=np.int32)
    assert np.array_equal(to_np(paddle.bitwise_xor(to_dev(xi), to_dev(yi))), ref.bitwise("xor", xi, yi))
    logits = (r.standard_normal((512, 8192)) * 2).astype(np.float32)
    lbl = r.integers(0, 8192, (512, 1), dtype=np.int64)
    sm, loss = ref.cross_entropy_with_softmax(logits, lbl)
    gl, gs = F.softmax_with_cross_entropy(to_dev(logits), to_dev(lbl), return_softmax=True)
    np.testing.assert_allclose(to_np(gs), sm, rtol=1e-5, atol=1e-30)
    np.testing.assert_allclose(to_np(gl), loss, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("shape", [(4, 300, 64), (2, 2051, 128), (3, 7, 24), (2, 48, 33)])
@pytest.mark.parametrize("dt", ["float32", "bfloat16"])
def test_softmax_strided_axis_special_values(port, shape, dt):
    """softmax along a middle axis (the vectorised column kernel and, for inner % V != 0, the scalar one): columns holding
    NaN, +inf, only -inf, a -1e4 outlier, or -inf in the rows visited first must follow the scalar rule of the CPU kernel."""
    r = _rng(12)
    x = (r.standard_normal(shape) * 2).astype(np.float32)
    x[0, 3, 1] = np.nan
    x[0, 5, 2] = np.inf
    x[0, :, 3] = -np.inf
    x[0, 2, 4] = -1e4
    x[0, : shape[1] // 2, 5] = -np.inf   # the running max stays -inf for the first half of the column
    x[1, 0, 6] = -np.inf
    if dt == "bfloat16":
        xb = port.f32_to_bf16(x)
        want = port.softmax(port.bf16_to_f32(xb), 1)
        got = port.bf16_to_f32(to_np(F.softmax(to_dev(xb, "bf16"), axis=1)))
        tol = dict(rtol=2 ** -7, atol=2e-28)
    else:
        want = port.softmax(x, 1)
        got = to_np(F.softmax(to_dev(x), axis=1))
        tol = dict(rtol=1e-5, atol=2e-28)
    for c in (1, 2, 3):
        assert np.isnan(want[0, :, c]).all() and np.isnan(got[0, :, c]).all(), f"column {c} must be NaN"
    mask = np.ones(shape, bool)
    mask[0, :, 1:4] = False
    np.testing.assert_allclose(got[mask], want[mask], **tol)
    if dt == "float32":
        lg = to_np(F.log_softmax(to_dev(x), axis=1))
        np.testing.assert_allclose(lg[mask], port.log_softmax(x, 1)[mask], rtol=1e-5, atol=1e-5)


def test_ref_argmin(ref, port):
    """argmin on the argmax engine (values mapped through an order-reversing bijection): every route, ties, special values."""
    r = _rng(13)
    x = r.integers(0, 50, (257, 4096), dtype=np.int64)  # many ties: lowest index must win
    for axis in (None, 0, 1, -1):
        assert np.array_equal(to_np(paddle.argmin(to_dev(x), axis=axis)), ref.argmin(x, axis)), axis
    xi = r.integers(-2**31, 2**31, (300, 5000), dtype=np.int32)
    xi[5, 7] = -2**31  # the value whose negation overflows: the mapping is bitwise NOT, not negation
    for axis in (0, 1):
        assert np.array_equal(to_np(paddle.argmin(to_dev(xi), axis=axis)), ref.argmin(xi, axis)), ("int32", axis)
    xf = r.standard_normal((64, 70000)).astype(np.float32)
    xf[1, 5] = np.nan
    xf[2, :] = np.inf
    xf[3, 100] = -np.inf
    xf[4, :] = np.nan
    assert np.array_equal(to_np(paddle.argmin(to_dev(xf), axis=1)), ref.argmin(xf, 1))
    assert np.array_equal(to_np(paddle.argmin(to_dev(xf), axis=1, dtype="int32")), ref.argmin(xf, 1, dtype=np.int32))
    for dt in (np.uint8, np.int16):
        info = np.iinfo(dt)
        big = r.integers(info.min, info.max, (3, 300, 4096), dtype=dt, endpoint=True)
        for axis in (0, 1, 2):
            assert np.array_equal(to_np(paddle.argmin(to_dev(big), axis=axis)), ref.argmin(big, axis)), (dt, axis)
    # 16-bit floats: the CPU kernel is not registered; the restatement applies the same rule to their float values
    for half, code, conv in (("bf16", O.BFLOAT16, port.f32_to_bf16), ("fp16", O.FLOAT16, port.f32_to_fp16)):
        xh = conv(r.standard_normal((33, 40000)).astype(np.float32))
        xh[0, :8] = conv(np.array([0.0, -0.0, 0.0, -0.0, 1, 1, 1, 1], np.float32))
        for axis in (None, 0, 1):
            assert np.array_equal(to_np(paddle.argmin(to_dev(xh, half), axis=axis)), port.argmin(xh, axis, code=code)), (half, axis)
        assert np.array_equal(to_np(paddle.argmax(to_dev(xh, half), axis=1)), port.argmax(xh, 1, code=code))
