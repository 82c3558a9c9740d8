"""GPU parity for log_softmax through the C-ABI: every route of the softmax engine (register rows,
TMA row pipeline, shared-memory rows, clusters, streaming, non-last axis) against the reference CPU kernel's recorded
outputs and the CPU oracle.  Bars: 1e-5 (relative, plus 1e-5 absolute for outputs near zero) in float32, 1e-12 in
float64, one unit in the last place for bf16 / fp16 outputs (computed in float, rounded once)."""
import os

import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402
from golden.make_golden_family import LOG_SOFTMAX_CASES  # noqa: E402

F = paddle.nn.functional
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_family_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLDEN))


def test_log_softmax_golden(gold):
    for name, _, axis in LOG_SOFTMAX_CASES:
        for dt in ("float32", "float64"):
            x = gold[f"lsm_{name}_{dt}__in0"]
            got = to_np(F.log_softmax(to_dev(x), axis=axis))
            tol = 1e-5 if dt == "float32" else 1e-12
            np.testing.assert_allclose(got, gold[f"lsm_{name}_{dt}__out"], rtol=tol, atol=tol)


ROUTES = [((1000, 64), -1), ((3000, 2048), -1), ((2500, 1024), -1), ((700, 1000), -1), ((64, 4096), -1), ((16, 40000), -1),
          ((6, 131072), -1), ((3, 600000), -1), ((5, 1001), -1), ((7, 300, 33), 1), ((9, 17, 128), 0), ((2, 50001), -1), ((1, 7), -1)]


@pytest.mark.parametrize("shape,axis", ROUTES)
def test_log_softmax_routes(port, shape, axis):
    rng = np.random.default_rng(21)
    x = (rng.standard_normal(shape) * 4).astype(np.float32)
    xd = x.astype(np.float64)
    # the float64 restatement is the yardstick: a float32 loop over 10^5 terms (the oracle's, and the CPU kernel's)
    # carries more summation error than the bar itself
    exact = port.log_softmax(xd, axis)
    np.testing.assert_allclose(to_np(F.log_softmax(to_dev(x), axis=axis)), exact, rtol=1e-5, atol=1e-5)
    if shape[axis] <= 4096:
        np.testing.assert_allclose(to_np(F.log_softmax(to_dev(x), axis=axis)), port.log_softmax(x, axis), rtol=1e-5, atol=1e-5)
    if np.prod(shape) <= 3_000_000:
        np.testing.assert_allclose(to_np(F.log_softmax(to_dev(xd), axis=axis)), exact, rtol=1e-12, atol=1e-12)
    for half, conv, code in (("bf16", port.f32_to_bf16, O.BFLOAT16), ("fp16", port.f32_to_fp16, O.FLOAT16)):
        xb = conv(x)
        got = to_np(F.log_softmax(to_dev(xb, half), axis=axis))
        xf = port.bf16_to_f32(xb) if half == "bf16" else port.fp16_to_f32(xb)
        want = conv(port.log_softmax(xf.astype(np.float64), axis).astype(np.float32))  # exact result, rounded once
        assert_half_close(got, want, 1, f"log_softmax {half} {shape}", half, atol=1e-5)


def test_log_softmax_edges(port):
    x = np.array([[1000.0, 0.0, -1000.0], [-np.inf, 0.0, 0.0], [5.0, 5.0, 5.0]], np.float32)
    np.testing.assert_allclose(to_np(F.log_softmax(to_dev(x))), port.log_softmax(x), rtol=1e-6, atol=1e-6)
    assert F.log_softmax(torch.tensor(3.0, device="cuda")).item() == 0.0
    assert tuple(F.log_softmax(torch.empty(0, 5, device="cuda")).shape) == (0, 5)
    with pytest.raises(ValueError, match="axis"):
        F.log_softmax(torch.zeros(3, 4, device="cuda"), axis=2)
    with pytest.raises(NotImplementedError):
        F.log_softmax(torch.zeros(3, 4, device="cuda", dtype=torch.int32))


def test_log_softmax_config4_sized():
    """config-4 shape (a quarter of it): exp(log_softmax) sums to one per row and equals softmax, with this library's own ops."""
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    x = (torch.randn(8, 32, 2048, 2048, device="cuda", generator=g) * 2).to(torch.bfloat16)
    ls = F.log_softmax(x, axis=-1)
    top = paddle.max(ls, axis=-1)
    assert float(paddle.max(top).item()) <= 0.0
    # the arg max of log_softmax is the arg max of x
    assert int(paddle.sum(paddle.not_equal(paddle.argmax(ls, axis=-1), paddle.argmax(x, axis=-1))).item()) <= x.numel() // 2048 // 100
    sm = F.softmax(x, axis=-1)
    assert int(paddle.sum(paddle.not_equal(paddle.argmax(sm, axis=-1), paddle.argmax(ls, axis=-1))).item()) <= x.numel() // 2048 // 100


@pytest.mark.parametrize("shape,axis", [((64, 2048), -1), ((33, 1001), -1), ((700, 4096), -1), ((1300, 512), -1), ((5, 40000), -1),
                                        ((8, 100, 64), 1), ((4, 17, 5, 3), 1), ((50, 1), -1), ((3, 7), 0)])
def test_log_softmax_grad_routes(port, shape, axis):
    """Backward of log_softmax (pdb200_log_softmax_grad, the softmax_grad kernels with the log rule): against the reference
    CPU kernel through the restatement pinned to it; 16-bit: <= 1 ulp of the rounded float result."""
    rng = np.random.default_rng(31)
    x = (rng.standard_normal(shape) * 2).astype(np.float32)
    y = port.log_softmax(x, axis)
    g = rng.standard_normal(shape).astype(np.float32)
    want = port.log_softmax_grad(y, g, axis)
    got = to_np(F.log_softmax_grad(to_dev(y), to_dev(g), axis=axis))
    scale = float(np.abs(g).sum(axis=axis).max())
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6 * (scale + 1))
    for half, conv, back in (("bf16", port.f32_to_bf16, port.bf16_to_f32), ("fp16", port.f32_to_fp16, port.fp16_to_f32)):
        yb, gb = conv(y), conv(g)
        wh = port.log_softmax_grad(back(yb), back(gb), axis)
        gh = to_np(F.log_softmax_grad(to_dev(yb, half), to_dev(gb, half), axis=axis))
        ulp = 2.0 ** -8 if half == "bf16" else 2.0 ** -11
        np.testing.assert_allclose(back(gh), wh, rtol=1.01 * ulp, atol=ulp * 1e-3 * (scale + 1) + 1e-6 * (scale + 1))
    xd = rng.standard_normal(shape)
    yd = xd - np.log(np.exp(xd).sum(axis=axis, keepdims=True))
    gd = rng.standard_normal(shape)
    np.testing.assert_allclose(to_np(F.log_softmax_grad(to_dev(yd), to_dev(gd), axis=axis)), port.log_softmax_grad(yd, gd, axis),
                               rtol=1e-12, atol=1e-12)


def test_log_softmax_grad_chain_and_edges(port):
    """Forward then backward against torch autograd in float64 on the same data; 0-D gives 0; in place on out_grad."""
    x = torch.randn(600, 4096, device="cuda")
    g = torch.randn(600, 4096, device="cuda")
    y = F.log_softmax(x, axis=-1)
    dx = F.log_softmax_grad(y, g, axis=-1)
    xr = x.double().requires_grad_(True)
    torch.log_softmax(xr, dim=-1).backward(g.double())
    np.testing.assert_allclose(to_np(dx), xr.grad.cpu().numpy(), rtol=1e-5, atol=1e-5)
    assert float(F.log_softmax_grad(torch.tensor(0.0, device="cuda"), torch.tensor(3.0, device="cuda"), axis=-1)) == 0.0
    with pytest.raises(ValueError):
        F.log_softmax_grad(y, g[:5], axis=-1)
