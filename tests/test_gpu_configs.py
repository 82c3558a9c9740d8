"""BASELINE.json's configurations at FULL size on the GPU.

The CPU oracle is single-threaded, so full tensors are checked (a) exactly against the oracle
on sampled rows / shards (every row of these ops is independent) and (b) through
size-independent properties: linearity of sums over shards ("checksum of checksums"), rows of a
softmax summing to 1, layer_norm rows having mean 0 / variance 1 before scale and bias,
idempotence of max.  Inputs come from seeded generators; nothing here reads /root/reference.
"""
import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional


def _gen(shape, dtype, seed, scale=1.0):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    return (torch.randn(shape, device="cuda", dtype=torch.float32, generator=g) * scale).to(dtype)


def test_c1_sum_fp32_4096(port):
    x = np.random.default_rng(1).random((4096, 4096), dtype=np.float32)
    got = to_np(paddle.sum(to_dev(x), axis=-1))
    np.testing.assert_allclose(got, port.reduce("sum", x, -1), rtol=1e-5)


def test_c2_add_bf16_and_int32_full(port):
    shape = (64, 2048, 4096)
    x = _gen(shape, torch.bfloat16, 2)
    y = _gen((4096,), torch.bfloat16, 22)
    out = paddle.add(x, y)
    assert out.shape == shape and out.dtype == torch.bfloat16
    xb, yb = to_np(x.view(-1, 4096)[[0, 1, 77777, 131071]]), to_np(y)
    want = port.elementwise("add", xb, yb, O.BFLOAT16)
    assert np.array_equal(to_np(out.view(-1, 4096)[[0, 1, 77777, 131071]]), want)
    # whole-tensor property: bit patterns equal to the same op done slab by slab (different
    # launch geometry, same arithmetic), compared on the device
    for s in range(0, 64, 16):
        part = paddle.add(x[s:s + 16], y)
        assert torch.equal(part.view(torch.int16), out[s:s + 16].view(torch.int16))
    del out, x
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    xi = torch.randint(-2**30, 2**30, shape, device="cuda", dtype=torch.int32, generator=g)
    yi = torch.randint(-2**30, 2**30, (4096,), device="cuda", dtype=torch.int32, generator=g)
    oi = paddle.add(xi, yi)
    rows = [0, 5, 131071]
    want = port.elementwise("add", to_np(xi.view(-1, 4096)[rows]), to_np(yi))
    assert np.array_equal(to_np(oi.view(-1, 4096)[rows]), want)
    # exact integer checksum: sum(out) == sum(x) + rows * sum(y)   (mod 2^64)
    sx, so, sy = paddle.sum(xi), paddle.sum(oi), paddle.sum(yi)
    assert (int(so) - int(sx) - 131072 * int(sy)) % (1 << 64) == 0


def test_c3_layer_norm_bf16_full(port):
    rows, D = 65536, 8192
    r = torch.arange(rows, device="cuda").view(-1, 1)
    x = (_gen((rows, D), torch.float32, 4) * (1 + r % 7) + r % 3).to(torch.bfloat16)
    g = torch.Generator(device="cuda"); g.manual_seed(44)
    w = (torch.rand(D, device="cuda", generator=g) + 0.5)
    b = (torch.rand(D, device="cuda", generator=g) - 0.5)
    for wp, bp in ((w.to(torch.bfloat16), b.to(torch.bfloat16)), (w, b)):
        y, m, v = F.layer_norm(x, [D], wp, bp, 1e-5, return_stats=True)
        idx = [0, 1, 2, 3, 6, 32767, 65534, 65535]
        xs = port.bf16_to_f32(to_np(x[idx]))
        wf = port.bf16_to_f32(to_np(wp)) if wp.dtype == torch.bfloat16 else to_np(wp)
        bf = port.bf16_to_f32(to_np(bp)) if bp.dtype == torch.bfloat16 else to_np(bp)
        oy, om, ov = port.layer_norm(xs, 1, wf, bf, 1e-5)
        # atol: see tests/test_gpu_parity.py::test_layer_norm_routes (the CPU kernel's sequential sums)
        assert_half_close(to_np(y[idx]), port.f32_to_bf16(oy), 1, "config 3 sampled rows", atol=2e-5)
        np.testing.assert_allclose(to_np(m[idx]), om, rtol=1e-5, atol=1e-5 * float(np.abs(xs).max()))
        np.testing.assert_allclose(to_np(v[idx]), ov, rtol=2e-5)
    # property over all rows: without scale/bias each row has mean ~0 and variance ~1
    y0 = F.layer_norm(x, [D])
    mu = paddle.mean(y0, axis=-1)
    assert float(mu.float().abs().max()) < 2e-2
    y0f = y0[:4096].float()
    assert torch.allclose((y0f * y0f).mean(-1), torch.ones(4096, device="cuda"), atol=2e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_c4_softmax_full(port, dtype):
    shape = (32, 32, 2048, 2048)
    x = _gen(shape, dtype, 5, scale=2.0)
    y = F.softmax(x, axis=-1)
    flat_x, flat_y = x.view(-1, 2048), y.view(-1, 2048)
    idx = [0, 1, 2047, 1048576, 2097151]
    if dtype == torch.float32:
        want = port.softmax(to_np(flat_x[idx]), -1)
        np.testing.assert_allclose(to_np(flat_y[idx]), want, rtol=1e-5)
    else:
        want = port.f32_to_bf16(port.softmax(port.bf16_to_f32(to_np(flat_x[idx])), -1))
        assert_half_close(to_np(flat_y[idx]), want, 1, "config 4 bf16 sampled rows")
    del x
    # property over every row: sums to 1
    s = paddle.sum(flat_y, axis=-1, dtype="float32") if dtype != torch.float32 else paddle.sum(flat_y, axis=-1)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert float((s.float() - 1).abs().max()) < tol


@pytest.mark.parametrize("dtype", [torch.int64, torch.float32])
def test_c5_full_tensor_reduce_one_shard(port, dtype):
    """One rank's shard of config 5: 2^31 elements (16 GiB int64 / 8 GiB fp32)."""
    n = 1 << 31
    g = torch.Generator(device="cuda"); g.manual_seed(6)
    if dtype == torch.int64:
        x = torch.randint(-2**40, 2**40, (n,), device="cuda", dtype=torch.int64, generator=g)
    else:
        x = torch.rand((n,), device="cuda", dtype=torch.float32, generator=g) * 2 - 1
    total = paddle.sum(x)
    mx = paddle.max(x)
    # checksum of checksums: 64 shards of 2^25, each also checked against the oracle for 2 of them
    parts = paddle.sum(x.view(64, -1), axis=1)
    pmax = paddle.max(x.view(64, -1), axis=1)
    for i in (0, 63):
        xs = to_np(x.view(64, -1)[i])
        if dtype == torch.int64:
            assert int(parts[i]) == int(port.reduce("sum", xs)) and int(pmax[i]) == int(xs.max())
        else:
            np.testing.assert_allclose(float(parts[i]), xs.astype(np.float64).sum(), rtol=0, atol=1e-5 * np.abs(xs).sum())
            assert float(pmax[i]) == float(xs.max())
    if dtype == torch.int64:
        assert int(total) == int(paddle.sum(parts))            # exact, wraps identically
        assert int(mx) == int(paddle.max(pmax))
    else:
        ref64 = float(parts.double().sum())
        assert abs(float(total) - ref64) <= 1e-5 * n * 0.5     # 1e-5 of sum |x| (E|x| = 0.5)
        assert float(mx) == float(paddle.max(pmax))            # max is exact and idempotent
    am = paddle.argmax(x)
    assert x[int(am)].item() == mx.item()
    assert int(am) == int((x == mx).nonzero()[0])              # lowest index among ties
