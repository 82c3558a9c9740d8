"""GPU parity for add_grad / multiply_grad through the C-ABI, against the reference CPU kernels'
recorded outputs and the CPU oracle.

Bars: integers bit-exact (wrapping sums are order-free); float32 / float64 to 1e-5 / 1e-12 of the sum of the terms'
magnitudes (the summation order differs from the CPU kernel's sequential loop — the reduction bar of BASELINE.json);
bf16 / fp16 within one unit in the last place of the result the reference's GPU arithmetic defines (products rounded
to the 16-bit type, summed in float, rounded once), or 1e-5 of the terms where the sum cancels.
"""
import os

import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, half_bits_to_f32, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402
from golden.make_golden_family import GRAD_CASES  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_family_golden.npz")


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


def _terms(op, x, y, d, which):
    """sum over the broadcast dims of |term| — the scale the float tolerance is relative to."""
    t = np.abs(d.astype(np.float64)) if op == "add" else np.abs(d.astype(np.float64) * (y if which == 0 else x).astype(np.float64))
    t = np.broadcast_to(t, d.shape)
    target = x.shape if which == 0 else y.shape
    lead = t.ndim - len(target)
    axes = tuple(range(lead)) + tuple(lead + i for i, s in enumerate(target) if s == 1 and t.shape[lead + i] != 1)
    return t.sum(axis=axes, keepdims=True).reshape(target) if axes else t.reshape(target)


def _check_float(op, x, y, d, got, want, rtol):
    for which, (g, w) in enumerate(zip(got, want)):
        if w is None:
            assert g is None
            continue
        tol = rtol * _terms(op, x, y, d, which) + np.finfo(x.dtype).tiny
        assert np.all(np.abs(to_np(g).astype(np.float64) - w.astype(np.float64)) <= tol), (op, which, x.shape, y.shape)


def test_grad_golden(gold):
    for name, _, _ in GRAD_CASES:
        for dt in ("float32", "float64", "int32", "int64"):
            x, y, d = (gold[f"grad_{name}_{dt}__in{i}"] for i in range(3))
            tx, ty, td = to_dev(x), to_dev(y), to_dev(d)
            for op, fn in (("add", paddle.add_grad), ("multiply", paddle.multiply_grad)):
                got = fn(tx, ty, td)
                want = (gold[f"grad_{op}_{name}_{dt}__out0"], gold[f"grad_{op}_{name}_{dt}__out1"])
                assert got[0].shape == tx.shape and got[1].shape == ty.shape
                if dt.startswith("int"):
                    assert np.array_equal(to_np(got[0]), want[0]) and np.array_equal(to_np(got[1]), want[1]), (op, name, dt)
                else:
                    _check_float(op, x, y, d, got, want, 1e-5 if dt == "float32" else 1e-12)


SHAPES = [((4096, 512), (512,)), ((512,), (4096, 512)), ((1000, 264), (264,)), ((70, 4096), (4096,)), ((3, 300, 256), (256,)),
          ((257, 1031), (1031,)), ((513, 2048), (513, 1)), ((7, 1, 33, 16), (5, 1, 16)), ((65, 70000), (65, 1)),
          ((64, 8, 128), (8, 1)), ((9, 10), ()), ((1 << 20,), (1 << 20,)), ((33, 1, 65), (7, 65)), ((2, 3, 4, 5), (1, 3, 1, 5))]


@pytest.mark.parametrize("xs,ys", SHAPES)
@pytest.mark.parametrize("want", [(True, True), (True, False), (False, True)])
def test_grad_routes(port, xs, ys, want):
    rng = np.random.default_rng(hash((xs, ys)) % 1000)
    os_ = np.broadcast_shapes(xs, ys)
    for dt in (np.float32, np.int64, np.int32, np.float64):
        if np.issubdtype(dt, np.integer):
            lim = 2**31 if dt == np.int64 else 2**15
            x, y, d = (rng.integers(-lim, lim, s_).astype(dt) for s_ in (xs, ys, os_))
        else:
            x, y, d = (rng.standard_normal(s_).astype(dt) for s_ in (xs, ys, os_))
        tx, ty, td = to_dev(x), to_dev(y), to_dev(d)
        for op, fn in (("add", paddle.add_grad), ("multiply", paddle.multiply_grad)):
            got = fn(tx, ty, td, want_x_grad=want[0], want_y_grad=want[1])
            ref = port.elementwise_grad(op, x, y, d, *want)
            if np.issubdtype(dt, np.integer):
                for g, w in zip(got, ref):
                    assert (g is None and w is None) or np.array_equal(to_np(g), w), (op, dt)
            else:
                _check_float(op, x, y, d, got, ref, 1e-5 if dt == np.float32 else 1e-12)


@pytest.mark.parametrize("xs,ys", SHAPES[:10])
@pytest.mark.parametrize("half,code", [("bf16", O.BFLOAT16), ("fp16", O.FLOAT16)])
def test_grad_16bit(port, xs, ys, half, code):
    rng = np.random.default_rng(3)
    conv = port.f32_to_bf16 if half == "bf16" else port.f32_to_fp16
    os_ = np.broadcast_shapes(xs, ys)
    xf, yf, df = (rng.standard_normal(s_).astype(np.float32) for s_ in (xs, ys, os_))
    x, y, d = conv(xf), conv(yf), conv(df)
    tx, ty, td = to_dev(x, half), to_dev(y, half), to_dev(d, half)
    for op, fn in (("add", paddle.add_grad), ("multiply", paddle.multiply_grad)):
        got = fn(tx, ty, td)
        want = port.elementwise_grad(op, x, y, d, code=code, gpu_arithmetic=True)
        xr, yr, dr = half_bits_to_f32(x, half), half_bits_to_f32(y, half), half_bits_to_f32(d, half)
        for which, (g, w) in enumerate(zip(got, want)):
            atol = float(1e-5 * _terms(op, xr, yr, dr, which).max())
            assert_half_close(to_np(g), w, 1, f"{op}_grad {half} d{'xy'[which]}", half, atol=atol)


def test_grad_in_place_and_errors():
    d = torch.randn(64, 128, device="cuda")
    x, y = torch.randn(64, 128, device="cuda"), torch.randn(128, device="cuda")
    dx, dy = paddle.add_grad(x, y, d)
    assert torch.equal(dx, d) and dx.data_ptr() != d.data_ptr()
    np.testing.assert_allclose(to_np(dy), to_np(d).astype(np.float64).sum(0), rtol=1e-5, atol=1e-4)
    with pytest.raises(ValueError, match="out_grad has shape"):
        paddle.add_grad(x, y, d[:32])
    with pytest.raises(ValueError, match="[Bb]roadcast"):
        paddle.multiply_grad(x, torch.randn(127, device="cuda"), d)
    with pytest.raises(NotImplementedError):
        paddle.multiply_grad(x.to(torch.int8), y.to(torch.int8), d.to(torch.int8))
    with pytest.raises(NotImplementedError, match="axis"):
        paddle.add_grad(x, y, d, axis=0)
    # a workspace smaller than asked for is refused, not overrun
    from paddlepaddle_b200 import _lib as L
    import ctypes
    lib = L.load()
    xd, xr = L.dims_array(x.shape)
    yd, yr = L.dims_array(y.shape)
    need = lib.pdb200_elementwise_grad_workspace_bytes(L.GRAD_MULTIPLY, xd, xr, yd, yr, L.FLOAT32, 0, 1)
    assert need > 0
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    out = torch.empty_like(y)
    rc = lib.pdb200_multiply_grad(x.data_ptr(), xd, xr, y.data_ptr(), yd, yr, d.data_ptr(), L.FLOAT32, None, out.data_ptr(),
                                  ws.data_ptr(), 16, None)
    assert rc == L.WORKSPACE_TOO_SMALL, lib.pdb200_last_error()


def test_config2_sized_multiply_grad():
    """BASELINE config 2 (x[64,2048,4096] bf16 * y[4096]): dx = dout*y must equal this library's own multiply;
    dy = column sums of round(dout*x) must equal the library's own multiply + sum pipeline to within the bar."""
    g = torch.Generator(device="cuda")
    g.manual_seed(4)
    x = torch.randn(64, 2048, 4096, device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randn(4096, device="cuda", generator=g).to(torch.bfloat16)
    d = torch.randn(64, 2048, 4096, device="cuda", generator=g).to(torch.bfloat16)
    dx, dy = paddle.multiply_grad(x, y, d)
    assert int(paddle.sum(paddle.not_equal(dx, paddle.multiply(d, y))).item()) == 0
    prod = paddle.multiply(d, x)
    want = paddle.sum(prod, axis=[0, 1], dtype="float32")
    mag = paddle.sum(paddle.multiply(prod, paddle.cast(paddle.greater_equal(prod, paddle.scale(prod, 0.0)), "bfloat16")),
                     axis=[0, 1], dtype="float32")  # sum of the positive terms: the scale of the cancellation
    got = paddle.cast(dy, "float32")
    err = paddle.subtract(got, want)
    ok = paddle.less_equal(paddle.maximum(err, paddle.scale(err, -1.0)),
                           paddle.add(paddle.scale(paddle.maximum(want, paddle.scale(want, -1.0)), 2.0**-8),
                                      paddle.scale(mag, 2e-5)))
    assert int(paddle.sum(ok).item()) == 4096
