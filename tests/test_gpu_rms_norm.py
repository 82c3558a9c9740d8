"""GPU parity for rms_norm through the C-ABI, both routes (TMA row pipeline, fallback), with and
without the fused bias + residual add.  The reference has no CPU kernel for this op, so the yardstick is the float64
restatement of the formula the reference's own test uses (parity unpinned against oracle/_ref — see the oracle's header).
Bars: 1e-5 relative (plus 1e-6 absolute) in float32; one unit in the last place of the 16-bit result for bf16 / fp16
(computed in float, rounded once), or 1e-5 of the terms where scale * x and the bias cancel."""
import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, half_bits_to_f32, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional

SHAPES = [((1000, 512), 1), ((700, 8192), 1), ((64, 9, 1024), 2), ((5, 1000), 1), ((3, 50000), 1), ((4, 6, 33), 1),
          ((600, 264), 1), ((2, 1), 1), ((0, 16), 1)]


@pytest.mark.parametrize("shape,axis", SHAPES)
@pytest.mark.parametrize("fused", [False, True])
def test_rms_norm_float32(port, shape, axis, fused):
    rng = np.random.default_rng(41)
    D = int(np.prod(shape[axis:]))
    x = (rng.standard_normal(shape) * 2 + 0.3).astype(np.float32)
    g, b = rng.uniform(0.5, 1.5, D).astype(np.float32), rng.uniform(-0.5, 0.5, D).astype(np.float32)
    res = rng.standard_normal(shape).astype(np.float32) if fused else None
    bias = rng.standard_normal(D).astype(np.float32) if fused else None
    for beta in (b, None):
        got = F.fused_rms_norm(to_dev(x).reshape(shape), to_dev(g), None if beta is None else to_dev(beta), 1e-5, axis,
                               bias=None if bias is None else to_dev(bias), residual=None if res is None else to_dev(res).reshape(shape))
        if x.size == 0:
            assert tuple(got[0].shape) == shape
            continue
        want, want_res, want_inv = port.rms_norm(x, g, beta, 1e-5, axis, bias=bias, residual=res)
        np.testing.assert_allclose(to_np(got[0]), want, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(to_np(got[-1]), want_inv, rtol=1e-5)
        if fused:
            np.testing.assert_allclose(to_np(got[1]), want_res, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("shape,axis", SHAPES[:7])
@pytest.mark.parametrize("half", ["bf16", "fp16"])
def test_rms_norm_16bit(port, shape, axis, half):
    rng = np.random.default_rng(42)
    conv = port.f32_to_bf16 if half == "bf16" else port.f32_to_fp16
    D = int(np.prod(shape[axis:]))
    x, res = conv((rng.standard_normal(shape) * 2).astype(np.float32)), conv(rng.standard_normal(shape).astype(np.float32))
    g, b = conv(rng.uniform(0.5, 1.5, D).astype(np.float32)), conv(rng.uniform(-0.5, 0.5, D).astype(np.float32))
    bias = conv(rng.standard_normal(D).astype(np.float32))
    f = lambda bits: half_bits_to_f32(bits, half).astype(np.float64)
    out, inv = F.fused_rms_norm(to_dev(x, half).reshape(shape), to_dev(g, half), to_dev(b, half), 1e-5, axis)
    want, _, want_inv = port.rms_norm(f(x), f(g), f(b), 1e-5, axis)
    assert_half_close(to_np(out), conv(want.astype(np.float32)), 1, f"rms_norm {half}", half, atol=1e-5 * float(np.abs(want).max() + 1))
    np.testing.assert_allclose(to_np(inv), want_inv, rtol=1e-5)
    out, ro, inv = F.fused_rms_norm(to_dev(x, half).reshape(shape), to_dev(g, half), None, 1e-5, axis, bias=to_dev(bias, half),
                                    residual=to_dev(res, half).reshape(shape))
    want, want_res, _ = port.rms_norm(f(x), f(g), None, 1e-5, axis, bias=f(bias), residual=f(res))
    assert_half_close(to_np(ro), conv(want_res.astype(np.float32)), 1, f"rms_norm residual_out {half}", half, atol=1e-6)
    # out is computed from the UNROUNDED sum (as the reference's ResidualAddBiasLoad hands it on)
    assert_half_close(to_np(out), conv(want.astype(np.float32)), 1, f"rms_norm fused {half}", half, atol=1e-5 * float(np.abs(want).max() + 1))


def test_rms_norm_in_place_residual_and_errors(port):
    rng = np.random.default_rng(43)
    x, res = rng.standard_normal((640, 1024)).astype(np.float32), rng.standard_normal((640, 1024)).astype(np.float32)
    g = rng.uniform(0.5, 1.5, 1024).astype(np.float32)
    from paddlepaddle_b200 import _lib as L
    import ctypes
    lib = L.load()
    tx, tr, tg = to_dev(x), to_dev(res), to_dev(g)
    out = torch.empty_like(tx)
    d, r = L.dims_array(tx.shape)
    rc = lib.pdb200_rms_norm(tx.data_ptr(), d, r, None, tr.data_ptr(), tg.data_ptr(), None, ctypes.c_float(1e-6), 1, L.FLOAT32,
                             out.data_ptr(), tr.data_ptr(), None, None)  # residual_out written over the residual
    assert rc == 0, lib.pdb200_last_error()
    want, want_res, _ = port.rms_norm(x, g, None, 1e-6, 1, residual=res)
    np.testing.assert_allclose(to_np(tr), want_res, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(to_np(out), want, rtol=1e-5, atol=1e-5)
    with pytest.raises(ValueError, match="bias is only added"):
        F.fused_rms_norm(tx, tg, bias=tg)
    with pytest.raises(ValueError, match="must have 1024 elements"):
        F.fused_rms_norm(tx, tg[:512])
    with pytest.raises(NotImplementedError):
        F.fused_rms_norm(tx.double(), tg.double())
    with pytest.raises(NotImplementedError, match="quantised"):
        F.fused_rms_norm(tx, tg, quant_scale=0.5)


def test_rms_norm_config3_sized():
    """config-3 shape (bf16 [65536, 8192]): every output row has unit root mean square when gamma = 1, checked with this
    library's own kernels; and the fused residual path equals add-then-normalise."""
    g = torch.Generator(device="cuda")
    g.manual_seed(6)
    x = (torch.randn(65536, 8192, device="cuda", generator=g) * 3).to(torch.bfloat16)
    ones = torch.ones(8192, device="cuda", dtype=torch.bfloat16)
    out, inv = F.fused_rms_norm(x, ones, None, 1e-6, 1)
    ms = paddle.mean(paddle.multiply(paddle.cast(out, "float32"), paddle.cast(out, "float32")), axis=-1)
    assert float(paddle.max(ms).item()) < 1.01 and float(paddle.min(ms).item()) > 0.99
    res = torch.randn(65536, 8192, device="cuda", generator=g).to(torch.bfloat16)
    out2, ro, _ = F.fused_rms_norm(x, ones, None, 1e-6, 1, residual=res)
    assert int(paddle.sum(paddle.not_equal(ro, paddle.add(x, res))).item()) == 0
