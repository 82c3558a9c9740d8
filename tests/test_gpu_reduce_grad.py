"""GPU parity for sum_grad / mean_grad through the C-ABI: bit-exact against the reference CPU
kernels' recorded outputs and the CPU oracle (a broadcast copy; mean divides by n exactly as the CPU kernel does)."""
import os

import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402
from golden.make_golden_family import REDUCE_GRAD_CASES  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "phi_family_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLDEN))


def _x(shape, dtype):
    return torch.empty(shape, dtype=dtype, device="cuda")


def test_reduce_grad_golden(gold):
    tdt = {"float32": torch.float32, "float64": torch.float64, "int32": torch.int32, "int64": torch.int64}
    for name, shape, axis, keep in REDUCE_GRAD_CASES:
        for dt in tdt:
            d = gold[f"rgrad_{name}_{dt}__in0"]
            td = to_dev(d).reshape(d.shape)
            for op, fn in (("sum", paddle.sum_grad), ("mean", paddle.mean_grad)):
                got = fn(_x(shape, tdt[dt]), td, axis=axis, keepdim=keep)
                assert tuple(got.shape) == tuple(shape)
                assert np.array_equal(to_np(got).reshape(shape), gold[f"rgrad_{op}_{name}_{dt}__out"]), (op, name, dt)


ROUTES = [((4096, 512), 0), ((4096, 512), 1), ((4096, 512), None), ((1000, 264), 0), ((257, 1031), 0), ((257, 1031), 1),
          ((64, 8, 128), 1), ((64, 8, 128), (0, 2)), ((7, 33, 5, 16), (1, 3)), ((1 << 20,), 0), ((3, 70000), 1), ((65, 1, 9), 1),
          ((2, 3, 4, 5, 6), (0, 2, 4)), ((16, 13, 2048), 1), ((3, 70, 1024), 1)]  # the last two: the batched inner route


@pytest.mark.parametrize("shape,axis", ROUTES)
@pytest.mark.parametrize("keep", [False, True])
def test_reduce_grad_routes(port, shape, axis, keep):
    rng = np.random.default_rng(7)
    axes = [] if axis is None else ([axis] if isinstance(axis, int) else list(axis))
    gshape = O.reduced_shape(shape, axes, keep)
    for dt, tdt in ((np.float32, torch.float32), (np.int64, torch.int64), (np.float64, torch.float64), (np.int32, torch.int32)):
        d = (rng.standard_normal(gshape) * 1000).astype(dt)
        td = to_dev(d).reshape(gshape)
        for op, fn in (("sum", paddle.sum_grad), ("mean", paddle.mean_grad)):
            got = fn(_x(shape, tdt), td, axis=axis, keepdim=keep)
            assert np.array_equal(to_np(got), port.reduce_grad(op, shape, d, axis, keep)), (op, dt)
    for half, conv, code, tdt in (("bf16", port.f32_to_bf16, O.BFLOAT16, torch.bfloat16), ("fp16", port.f32_to_fp16, O.FLOAT16, torch.float16)):
        d = conv(rng.standard_normal(gshape).astype(np.float32))
        td = to_dev(d, half).reshape(gshape)
        for op, fn in (("sum", paddle.sum_grad), ("mean", paddle.mean_grad)):
            got = fn(_x(shape, tdt), td, axis=axis, keepdim=keep)
            assert np.array_equal(to_np(got), port.reduce_grad(op, shape, d, axis, keep, code=code)), (op, half)


def test_reduce_grad_unaligned_and_errors():
    g = torch.randn(1031 + 3, device="cuda")[3:]  # not 16-byte aligned
    got = paddle.sum_grad(_x((77, 1031), torch.float32), g, axis=0)
    assert torch.equal(got, g.expand(77, 1031))
    with pytest.raises(ValueError, match="out_grad has shape"):
        paddle.sum_grad(_x((4, 5), torch.float32), torch.zeros(4, device="cuda"), axis=0)
    with pytest.raises(ValueError):
        paddle.mean_grad(_x((4, 5), torch.float32), torch.zeros(4, device="cuda"), axis=2)
    with pytest.raises(NotImplementedError, match="not registered"):
        paddle.sum_grad(_x((4, 5), torch.int8), torch.zeros(5, device="cuda", dtype=torch.int8), axis=0)


def test_round_trip_with_forward_at_config_sizes():
    """sum(sum_grad(g)) over the same axes is n * g; checked with this library's own forward at a config-3-sized tensor."""
    g = torch.randn(8192, device="cuda")
    xg = paddle.sum_grad(_x((65536, 8192), torch.float32), g, axis=0)          # 2 GiB written
    back = paddle.sum(xg, axis=0)
    np.testing.assert_allclose(to_np(back), to_np(g).astype(np.float64) * 65536, rtol=1e-5)  # the fp32 reduction bar
    gm = torch.randn(65536, device="cuda").to(torch.bfloat16)
    xm = paddle.mean_grad(_x((65536, 8192), torch.bfloat16), gm, axis=-1)
    assert int(paddle.sum(paddle.not_equal(paddle.max(xm, axis=-1), paddle.min(xm, axis=-1))).item()) == 0  # constant rows
    want = paddle.cast(paddle.scale(paddle.cast(gm, "float32"), 1.0 / 8192), "bfloat16")  # /8192 is exact in binary
    assert int(paddle.sum(paddle.not_equal(paddle.max(xm, axis=-1), want)).item()) == 0
