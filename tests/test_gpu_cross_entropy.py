"""GPU parity for cross_entropy_with_softmax through the C-ABI, against the reference CPU kernel's
recorded outputs and the CPU oracle: softmax to the softmax bars (1e-5 float32, 1e-12 float64), loss to 1e-5 / 1e-12
(relative, plus the same absolute term for losses near zero); float16 / bfloat16: the loss is computed in float from the
logits and the row statistics and rounded once — one unit in the last place of the 16-bit result."""
import os

import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402
from golden.make_golden_family import CE_CASES  # noqa: E402

F = paddle.nn.functional
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_family_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLDEN))


def test_cross_entropy_golden(gold):
    for name, shape, axis in CE_CASES:
        for dt in ("float32", "float64"):
            x, hard, soft = (gold[f"ce_{name}_{dt}__in{i}"] for i in range(3))
            tol = dict(rtol=1e-5, atol=1e-5) if dt == "float32" else dict(rtol=1e-12, atol=1e-12)
            for kind, lbl in (("hard", hard), ("soft", soft)):
                loss, sm = F.softmax_with_cross_entropy(to_dev(x), to_dev(lbl).reshape(lbl.shape), soft_label=kind == "soft",
                                                        axis=axis, return_softmax=True)
                np.testing.assert_allclose(to_np(sm), gold[f"ce_{kind}_{name}_{dt}__out0"], **tol)
                np.testing.assert_allclose(to_np(loss), gold[f"ce_{kind}_{name}_{dt}__out1"], **tol)


@pytest.mark.parametrize("shape,axis", [((4096, 512), -1), ((300, 32000), -1), ((64, 7, 1000), 1), ((2, 128256), -1), ((33, 5, 17), 0),
                                        ((1000, 1), -1)])
def test_cross_entropy_routes(port, shape, axis):
    rng = np.random.default_rng(31)
    ax = axis % len(shape)
    lshape = tuple(1 if i == ax else d for i, d in enumerate(shape))
    x = (rng.standard_normal(shape) * 3).astype(np.float32)
    hard = rng.integers(0, shape[ax], lshape)
    hard.flat[::7] = -1  # ignored positions
    soft = rng.random(shape).astype(np.float32)
    soft /= soft.sum(axis=ax, keepdims=True)
    # yardstick: the float64 restatement (long float32 sums carry more error than the bar)
    sm64, lh64 = port.cross_entropy_with_softmax(x.astype(np.float64), hard, False, True, -1, axis)
    _, ls64 = port.cross_entropy_with_softmax(x.astype(np.float64), soft.astype(np.float64), True, True, -1, axis)
    for lbl_t in (torch.int64, torch.int32):
        loss, sm = F.softmax_with_cross_entropy(to_dev(x), to_dev(hard).reshape(lshape).to(lbl_t), ignore_index=-1, axis=axis,
                                                return_softmax=True)
        np.testing.assert_allclose(to_np(sm), sm64, rtol=1e-5, atol=1e-9)
        np.testing.assert_allclose(to_np(loss), lh64, rtol=1e-5, atol=1e-5)
    loss = F.softmax_with_cross_entropy(to_dev(x), to_dev(soft), soft_label=True, axis=axis)
    np.testing.assert_allclose(to_np(loss), ls64, rtol=1e-5, atol=1e-5)
    xd = x.astype(np.float64)
    loss = F.softmax_with_cross_entropy(to_dev(xd), to_dev(hard).reshape(lshape), ignore_index=-1, axis=axis)
    np.testing.assert_allclose(to_np(loss), lh64, rtol=1e-12, atol=1e-12)
    for half, conv, back in (("fp16", port.f32_to_fp16, port.fp16_to_f32), ("bf16", port.f32_to_bf16, port.bf16_to_f32)):
        xh = conv(x)
        xr = back(xh).astype(np.float64)  # the values the kernel sees
        _, want_h = port.cross_entropy_with_softmax(xr, hard, False, True, -1, axis)
        _, want_s = port.cross_entropy_with_softmax(xr, soft.astype(np.float64), True, True, -1, axis)
        got_h = F.softmax_with_cross_entropy(to_dev(xh, half), to_dev(hard).reshape(lshape), ignore_index=-1, axis=axis)
        assert_half_close(to_np(got_h), conv(want_h.astype(np.float32)), 1, f"ce hard {half}", half, atol=1e-5)
        sl = conv(soft)
        _, want_s = port.cross_entropy_with_softmax(xr, back(sl).astype(np.float64), True, True, -1, axis)
        got_s = F.softmax_with_cross_entropy(to_dev(xh, half), to_dev(sl, half), soft_label=True, axis=axis)
        assert_half_close(to_np(got_s), conv(want_s.astype(np.float32)), 1, f"ce soft {half}", half, atol=1e-4)


def test_cross_entropy_input_is_softmax_and_errors(port):
    rng = np.random.default_rng(32)
    x = (rng.standard_normal((50, 33)) * 2).astype(np.float32)
    prob = port.softmax(x, -1)
    hard = rng.integers(0, 33, (50, 1))
    lib_loss, lib_sm = F._cross_entropy(to_dev(prob), to_dev(hard).reshape(50, 1), False, False, -100, -1, True)
    want_sm, want_loss = port.cross_entropy_with_softmax(prob, hard, False, False, -100, -1)
    assert np.array_equal(to_np(lib_sm), want_sm)
    np.testing.assert_allclose(to_np(lib_loss), want_loss, rtol=1e-6, atol=1e-6)
    bad = hard.copy()
    bad[3, 0] = 33  # out of range: the reference raises; here the position is NaN and the rest is unaffected
    loss = to_np(F.softmax_with_cross_entropy(to_dev(x), to_dev(bad).reshape(50, 1)))
    assert np.isnan(loss[3, 0]) and not np.isnan(np.delete(loss, 3, 0)).any()
    with pytest.raises(ValueError, match="label has shape"):
        F.softmax_with_cross_entropy(to_dev(x), to_dev(hard).reshape(50))
    with pytest.raises(TypeError):
        F.softmax_with_cross_entropy(to_dev(x), to_dev(hard.astype(np.float32)).reshape(50, 1))
    with pytest.raises(NotImplementedError):
        F.softmax_with_cross_entropy(to_dev(x).to(torch.int32), to_dev(hard).reshape(50, 1))
    # in place (softmax written over the logits): the loss falls back to -log(softmax) and still meets the float32 bar
    from paddlepaddle_b200 import _lib as L
    import ctypes
    lib = L.load()
    tx = to_dev(x)
    tl = to_dev(hard).reshape(50, 1)
    loss = torch.empty(50, 1, device="cuda")
    d, r = L.dims_array(tx.shape)
    rc = lib.pdb200_cross_entropy_with_softmax(tx.data_ptr(), d, r, tl.data_ptr(), L.INT64, 0, 1, -100, -1, L.FLOAT32, tx.data_ptr(),
                                               loss.data_ptr(), None, 0, None)
    assert rc == 0, lib.pdb200_last_error()
    want_sm, want_loss = port.cross_entropy_with_softmax(x, hard, False, True, -100, -1)
    np.testing.assert_allclose(to_np(tx), want_sm, rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(to_np(loss), want_loss, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("shape,axis", [((700, 4096), -1), ((64, 1000), 1), ((33, 1001), -1), ((4, 9, 5), 1), ((3, 5, 4, 6), 2),
                                        ((2, 50000), -1), ((5, 1), -1)])
def test_cross_entropy_grad_against_compiled_reference(ref, port, shape, axis):
    """logits_grad (pdb200_cross_entropy_with_softmax_grad) bit-exact against the reference CPU kernel for float32 / float64,
    hard labels (int64 and int32, with ignored positions) and soft labels; float16 / bfloat16 within one ulp of the rounded
    float result."""
    rng = np.random.default_rng(9)
    ax = axis % len(shape)
    lshape = tuple(1 if i == ax else d for i, d in enumerate(shape))
    for dt in (np.float32, np.float64):
        x = (rng.standard_normal(shape) * 2).astype(dt)
        sm = port.softmax(x, axis)
        g = rng.standard_normal(lshape).astype(dt)
        lbl = rng.integers(0, shape[ax], lshape).astype(np.int64)
        lbl.reshape(-1)[::3] = -100
        want = ref.cross_entropy_with_softmax_grad(lbl, sm, g, False, -100, axis)
        for ldt in (np.int64, np.int32):
            got = F.softmax_with_cross_entropy_grad(to_dev(lbl.astype(ldt)), to_dev(sm), to_dev(g), soft_label=False, ignore_index=-100, axis=axis)
            assert np.array_equal(to_np(got), want), (dt, ldt)
        y = rng.dirichlet(np.ones(shape[ax]), size=int(np.prod(lshape))).astype(dt)
        y = np.moveaxis(y.reshape(tuple(d for i, d in enumerate(shape) if i != ax) + (shape[ax],)), -1, ax).copy()
        want = ref.cross_entropy_with_softmax_grad(y, sm, g, True, -100, axis)
        got = F.softmax_with_cross_entropy_grad(to_dev(y), to_dev(sm), to_dev(g), soft_label=True, axis=axis)
        assert np.array_equal(to_np(got), want), (dt, "soft")
    x = (rng.standard_normal(shape) * 2).astype(np.float32)
    sm = port.softmax(x, axis)
    g = rng.standard_normal(lshape).astype(np.float32)
    lbl = rng.integers(0, shape[ax], lshape).astype(np.int64)
    for half, conv, back in (("bf16", port.f32_to_bf16, port.bf16_to_f32), ("fp16", port.f32_to_fp16, port.fp16_to_f32)):
        sb, gb = conv(sm), conv(g)
        want = port.cross_entropy_with_softmax_grad(lbl, back(sb), back(gb), False, -100, axis)
        got = to_np(F.softmax_with_cross_entropy_grad(to_dev(lbl), to_dev(sb, half), to_dev(gb, half), axis=axis))
        assert_half_close(got, conv(want), 1, f"ce grad {half} {shape}", kind=half, atol=1e-6)


def test_cross_entropy_forward_backward_chain():
    """softmax_with_cross_entropy then its backward against torch autograd in float64 (mean-reduced loss), config-sized rows."""
    n, c = 4096, 8192
    x = torch.randn(n, c, device="cuda")
    lbl = torch.randint(0, c, (n, 1), device="cuda")
    loss, sm = F.softmax_with_cross_entropy(x, lbl, return_softmax=True)
    g = torch.full((n, 1), 1.0 / n, device="cuda")
    dx = F.softmax_with_cross_entropy_grad(lbl, sm, g)
    xr = x.double().requires_grad_(True)
    torch.nn.functional.cross_entropy(xr, lbl.view(-1)).backward()
    np.testing.assert_allclose(to_np(dx), xr.grad.cpu().numpy(), rtol=1e-4, atol=1e-9)
    with pytest.raises(ValueError):
        F.softmax_with_cross_entropy_grad(lbl, sm, g[:5])
