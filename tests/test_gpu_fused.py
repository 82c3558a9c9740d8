"""GPU parity of the fused neighbours of softmax / layer_norm: masked softmax (additive mask
shared by the heads; causal upper triangle) and layer_norm(residual + dropout(x + bias)).  The reference kernels are
GPU-only; the oracle pins the math THROUGH the reference CPU softmax / layer_norm kernels on pre-combined inputs
(tests/oracle_lib.py).  Bars: fp32 rtol 1e-5 (atol 2e-28 covers the CPU kernel's exp(-64) clip, which the GPU kernels
of the reference do not have); 1 ulp of the output type for 16-bit outputs; zeros above the diagonal exact."""
import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional


@pytest.fixture(scope="module")
def cpu(port):
    return O.Oracle("ref") if O.have_ref() else port


def _half(port, a, dt):
    if dt == "bfloat16":
        return port.f32_to_bf16(a), "bf16", port.bf16_to_f32, port.f32_to_bf16
    if dt == "float16":
        return port.f32_to_fp16(a), "fp16", port.fp16_to_f32, port.f32_to_fp16
    return a, None, (lambda v: v), (lambda v: v)


@pytest.mark.parametrize("shape", [(2, 4, 128, 128), (1, 3, 512, 2048), (2, 2, 300, 1024), (1, 3, 40, 100), (1, 2, 8, 8192)])
@pytest.mark.parametrize("dt,mask_dt", [("float32", "float32"), ("bfloat16", "bfloat16"), ("float16", "float32"),
                                        ("float16", "float16")])
def test_softmax_mask_fuse(port, cpu, shape, dt, mask_dt):
    rng = np.random.default_rng(hash((shape, dt, mask_dt)) & 0xffff)
    B, H, Sq, Sk = shape
    x = (rng.standard_normal(shape) * 2).astype(np.float32)
    mask = np.where(rng.random((B, 1, Sq, Sk)) < 0.3, np.float32(-10000.0), np.float32(0.0)).astype(np.float32)
    mask[:, :, :, 0] = 0.0  # no row fully masked
    mask[0, 0, 1, 1:3] = -np.inf
    mask += (rng.standard_normal(mask.shape) * 0.1).astype(np.float32) * (mask == 0)
    xb, xh, up, down = _half(port, x, dt)
    mb, mh, mup, _ = _half(port, mask, mask_dt)
    want = cpu.softmax_mask_fuse(up(xb), mup(mb))
    got = F.softmax_mask_fuse(to_dev(xb, xh), to_dev(mb, mh))
    assert tuple(got.shape) == shape
    if xh is None:
        np.testing.assert_allclose(to_np(got), want, rtol=1e-5, atol=2e-28)
    else:
        assert_half_close(to_np(got), down(want), 1, f"softmax_mask_fuse {shape} {dt}", kind=xh, atol=2e-28)


def test_softmax_mask_fuse_errors():
    x = torch.zeros((2, 2, 4, 64), device="cuda")
    with pytest.raises(ValueError, match="second dim must be 1"):
        F.softmax_mask_fuse(x, torch.zeros((2, 2, 4, 64), device="cuda"))
    with pytest.raises(ValueError, match="should be equal"):
        F.softmax_mask_fuse(x, torch.zeros((2, 1, 4, 32), device="cuda"))
    with pytest.raises(TypeError):
        F.softmax_mask_fuse(x, torch.zeros((2, 1, 4, 64), device="cuda", dtype=torch.float16))


@pytest.mark.parametrize("shape", [(2, 4, 128, 128), (1, 2, 2048, 2048), (3, 2, 304, 304), (1, 2, 100, 100), (2, 150, 32, 32)])
@pytest.mark.parametrize("dt", ["float32", "bfloat16", "float16"])
def test_softmax_mask_fuse_upper_triangle(port, cpu, shape, dt):
    rng = np.random.default_rng(hash((shape, dt, "tri")) & 0xffff)
    x = (rng.standard_normal(shape) * 2).astype(np.float32)
    x[0, 0, 5, 2] = -np.inf      # a masked-out score below the diagonal
    x[0, 1, 3, 20:] = np.nan     # garbage ABOVE the diagonal must never be read into the result
    xb, xh, up, down = _half(port, x, dt)
    want = cpu.softmax_mask_fuse_upper_triangle(up(xb))  # the oracle reads row prefixes only
    got = to_np(F.softmax_mask_fuse_upper_triangle(to_dev(xb, xh)))
    S = shape[-1]
    upper = np.triu(np.ones((S, S), bool), 1)
    gz = got[..., upper]
    assert not gz.any(), "entries above the diagonal must be exactly zero"
    if xh is None:
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=2e-28)
    else:
        assert_half_close(got, down(want), 1, f"upper_triangle {shape} {dt}", kind=xh, atol=2e-28)


def test_softmax_mask_fuse_upper_triangle_errors():
    with pytest.raises(ValueError, match="Key seq len must be equal"):
        F.softmax_mask_fuse_upper_triangle(torch.zeros((1, 1, 32, 64), device="cuda"))


@pytest.mark.parametrize("rows,D,dt,pdt,has_bias,training,mode,p", [
    (4096, 8192, "bfloat16", "bfloat16", True, True, "upscale_in_train", 0.0),
    (2048, 1024, "float32", "float32", False, True, "upscale_in_train", 0.0),
    (1024, 4096, "float16", "float32", True, False, "downscale_in_infer", 0.3),
    (1000, 2048, "bfloat16", "float32", True, False, "upscale_in_train", 0.5),
    (700, 768, "float32", "float32", True, False, "downscale_in_infer", 0.25),
    (7, 100, "float32", "float32", True, True, "upscale_in_train", 0.0),
    (33, 1000, "float16", "float16", False, False, "downscale_in_infer", 0.1),
])
def test_fused_bias_dropout_residual_layer_norm(port, cpu, rows, D, dt, pdt, has_bias, training, mode, p):
    rng = np.random.default_rng(hash((rows, D, dt, pdt)) & 0xffff)
    x = (rng.standard_normal((rows, D)) * 2).astype(np.float32)
    res = (rng.standard_normal((rows, D)) * 2 + 0.5).astype(np.float32)
    bias = rng.uniform(-1, 1, D).astype(np.float32) if has_bias else None
    w = rng.uniform(0.5, 1.5, D).astype(np.float32)
    b = rng.uniform(-0.5, 0.5, D).astype(np.float32)
    xb, xh, up, down = _half(port, x, dt)
    rb = _half(port, res, dt)[0]
    bb = _half(port, bias, dt)[0] if has_bias else None
    wb, wh, wup, _ = _half(port, w, pdt)
    lb = _half(port, b, pdt)[0]
    factor = 1.0 if training or mode == "upscale_in_train" else 1.0 - p
    wy, wd, wm, wv = cpu.fused_bias_dropout_residual_layer_norm(xb, rb, bb, wup(wb), wup(lb), 1e-5, factor, half=xh, conv=port)
    y, d, m, v = F.fused_bias_dropout_residual_layer_norm(
        to_dev(xb, xh), to_dev(rb, xh), to_dev(bb, xh) if has_bias else None, to_dev(wb, wh), to_dev(lb, wh),
        dropout_rate=p, ln_epsilon=1e-5, training=training, mode=mode, return_all=True)
    if xh is None:
        assert np.array_equal(to_np(d), wd), "bias_dropout_residual_out must be bit-exact"
        np.testing.assert_allclose(to_np(y), wy, rtol=1e-5, atol=2e-5)
    else:
        assert np.array_equal(to_np(d), wd), "bias_dropout_residual_out must be bit-exact (T arithmetic, step by step)"
        assert_half_close(to_np(y), down(wy), 1, f"fused ln y [{rows},{D}] {dt}", kind=xh, atol=2e-5)
    np.testing.assert_allclose(to_np(m), wm, rtol=1e-5, atol=1e-5 * float(np.abs(up(wd)).max()))
    np.testing.assert_allclose(to_np(v), wv, rtol=2e-5)


def test_fused_bias_dropout_residual_layer_norm_float64_and_errors(cpu):
    rng = np.random.default_rng(11)
    x, res = rng.standard_normal((16, 64)), rng.standard_normal((16, 64))
    w, b = rng.uniform(0.5, 1.5, 64), rng.uniform(-0.5, 0.5, 64)
    wy, wm, wv = cpu.layer_norm(x + res, 1, w, b, 1e-5)
    y, d, m, v = F.fused_bias_dropout_residual_layer_norm(to_dev(x), to_dev(res), None, to_dev(w), to_dev(b), dropout_rate=0.0,
                                                          return_all=True)
    np.testing.assert_allclose(to_np(y), wy, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(to_np(m), wm, rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(to_np(v), wv, rtol=1e-12)
    with pytest.raises(NotImplementedError):
        F.fused_bias_dropout_residual_layer_norm(to_dev(x), to_dev(res), dropout_rate=0.5, training=True)
    with pytest.raises(ValueError):
        F.fused_bias_dropout_residual_layer_norm(to_dev(x), to_dev(res), mode="nope", dropout_rate=0.0)


@pytest.mark.parametrize("rows,D,dt,has_bias,alpha", [
    (4096, 8192, "bfloat16", True, 1.0), (2048, 1024, "float32", False, 0.5), (1024, 4096, "float16", True, 0.7071),
    (1000, 2048, "bfloat16", True, 2.0), (700, 768, "float32", True, 1.5), (7, 100, "float32", True, 0.3),
    (33, 1000, "float16", False, 1.0),
])
def test_fused_layer_norm_with_residual(port, cpu, rows, D, dt, has_bias, alpha):
    """fused_bias_residual_layernorm, the forms with a residual: layer_norm(x + bias + alpha * residual) with float32
    parameters, and the sum alone when no norm parameter is given.  The sum is bit-exact (T arithmetic step by step, as
    the reference's helper), the normalisation goes through the reference CPU layer_norm on float(sum)."""
    rng = np.random.default_rng(hash((rows, D, dt, alpha)) & 0xffff)
    x = (rng.standard_normal((rows, D)) * 2).astype(np.float32)
    res = (rng.standard_normal((rows, D)) * 2 + 0.5).astype(np.float32)
    bias = rng.uniform(-1, 1, D).astype(np.float32) if has_bias else None
    w = rng.uniform(0.5, 1.5, D).astype(np.float32)
    b = rng.uniform(-0.5, 0.5, D).astype(np.float32)
    xb, xh, up, down = _half(port, x, dt)
    rb = _half(port, res, dt)[0]
    bb = _half(port, bias, dt)[0] if has_bias else None
    wy, wd, wm, wv = cpu.fused_bias_dropout_residual_layer_norm(xb, rb, bb, w, b, 1e-5, 1.0, half=xh, conv=port, alpha=alpha)
    dx, dr, db = to_dev(xb, xh), to_dev(rb, xh), (to_dev(bb, xh) if has_bias else None)
    y, d, m, v = F.fused_layer_norm(dx, to_dev(w), to_dev(b), 1e-5, residual_alpha=alpha, begin_norm_axis=1, bias=db, residual=dr,
                                    return_all=True)
    assert np.array_equal(to_np(d), wd), "residual_out must be bit-exact"
    if xh is None:
        np.testing.assert_allclose(to_np(y), wy, rtol=1e-5, atol=2e-5)
    else:
        assert_half_close(to_np(y), down(wy), 1, f"fused_layer_norm y [{rows},{D}] {dt}", kind=xh, atol=2e-5)
    np.testing.assert_allclose(to_np(m), wm, rtol=1e-5, atol=1e-5 * float(np.abs(up(wd)).max()))
    np.testing.assert_allclose(to_np(v), wv, rtol=2e-5)
    if xh is None and cpu.kind == "ref" and cpu.have_avx512():
        # the reference's own CPU kernel of this op (AVX-512; it adds the residual before the bias and uses E[d^2] - mean^2)
        co, cro, cm, crstd = cpu.fused_layer_norm_cpu(x, bias, res, w, b, 1e-5, alpha)
        np.testing.assert_allclose(to_np(d), cro, rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(to_np(y), co, rtol=1e-4, atol=1e-4)
        np.testing.assert_allclose(1.0 / np.sqrt(to_np(v) + 1e-5), crstd, rtol=5e-5)
    # the two-output form the Python API returns, and only one norm parameter
    y2, d2 = F.fused_layer_norm(dx, to_dev(w), None, 1e-5, residual_alpha=alpha, bias=db, residual=dr)
    assert np.array_equal(to_np(d2), wd)
    wy2 = cpu.fused_bias_dropout_residual_layer_norm(xb, rb, bb, w, None, 1e-5, 1.0, half=xh, conv=port, alpha=alpha)[0]
    if xh is None:
        np.testing.assert_allclose(to_np(y2), wy2, rtol=1e-5, atol=2e-5)
    else:
        assert_half_close(to_np(y2), down(wy2), 1, "fused_layer_norm, weight only", kind=xh, atol=2e-5)
    # no norm parameters: the fused sum alone
    s = F.fused_layer_norm(dx, None, None, 1e-5, residual_alpha=alpha, bias=db, residual=dr)
    assert np.array_equal(to_np(s), cpu.fused_bias_dropout_residual_layer_norm(xb, rb, bb, None, None, 1e-5, 1.0, half=xh, conv=port,
                                                                             alpha=alpha, sum_only=True))


def test_fused_layer_norm_plain_axes_and_errors(port, cpu):
    """Without a residual the op is layer_norm(x) with float32 parameters (bias is not read); begin_norm_axis folds the
    trailing dimensions; an unaligned / ragged row takes the one-CTA-per-row route."""
    rng = np.random.default_rng(5)
    x = (rng.standard_normal((6, 50, 40)) * 3 + 1).astype(np.float32)
    w, b = rng.uniform(0.5, 1.5, 2000).astype(np.float32), rng.uniform(-0.5, 0.5, 2000).astype(np.float32)
    wy, wm, wv = cpu.layer_norm(x.reshape(6, 2000), 1, w, b, 1e-6)
    y, r, m, v = F.fused_layer_norm(to_dev(x), to_dev(w), to_dev(b), 1e-6, begin_norm_axis=1, return_all=True)
    assert r is None
    np.testing.assert_allclose(to_np(y).reshape(6, 2000), wy, rtol=1e-5, atol=2e-5)
    np.testing.assert_allclose(to_np(m), wm, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(to_np(v), wv, rtol=2e-5)
    bits = port.f32_to_bf16(x)
    yb = F.fused_layer_norm(to_dev(bits, "bf16"), to_dev(w), to_dev(b), 1e-6, begin_norm_axis=1,
                            bias=to_dev(port.f32_to_bf16(b), "bf16"))  # bias ignored on this branch, as in the reference
    wyb = cpu.layer_norm(port.bf16_to_f32(bits).reshape(6, 2000), 1, w, b, 1e-6)[0]
    assert_half_close(to_np(yb).reshape(6, 2000), port.f32_to_bf16(wyb), 1, "fused_layer_norm plain bf16", kind="bf16", atol=2e-5)
    # residual + begin_norm_axis = 2 + a ragged row length
    xr = (rng.standard_normal((5, 7, 33)) * 2).astype(np.float32)
    rr = (rng.standard_normal((5, 7, 33)) * 2).astype(np.float32)
    w3, b3 = rng.uniform(0.5, 1.5, 33).astype(np.float32), rng.uniform(-0.5, 0.5, 33).astype(np.float32)
    wy, wd, wm, wv = cpu.fused_bias_dropout_residual_layer_norm(xr.reshape(35, 33), rr.reshape(35, 33), None, w3, b3, 1e-5, 1.0, alpha=0.25)
    y, d, m, v = F.fused_layer_norm(to_dev(xr), to_dev(w3), to_dev(b3), 1e-5, residual_alpha=0.25, begin_norm_axis=2,
                                    residual=to_dev(rr), return_all=True)
    assert np.array_equal(to_np(d).reshape(35, 33), wd)
    np.testing.assert_allclose(to_np(y).reshape(35, 33), wy, rtol=1e-5, atol=2e-5)
    np.testing.assert_allclose(to_np(v), wv, rtol=2e-5)
    with pytest.raises(NotImplementedError):
        F.fused_layer_norm(to_dev(xr), to_dev(w3), to_dev(b3), 1e-5, begin_norm_axis=2, quant_scale=0.5)
    with pytest.raises(ValueError):
        F.fused_layer_norm(to_dev(xr), to_dev(w3.astype(np.float64)), None, 1e-5, begin_norm_axis=2)
    with pytest.raises(ValueError):
        F.fused_layer_norm(to_dev(xr), to_dev(w3), None, 1e-5, begin_norm_axis=2, residual=to_dev(rr[:2]))
