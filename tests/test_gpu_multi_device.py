"""One process driving two GPUs: per-device state inside libpdb200.so (function attributes for > 48 KB of dynamic shared
memory, cached launch geometry, the built-in scratch of pdb200_sum_axis0_static) must be kept per device.  Skipped on a
single-GPU box; run with `gpurun --gpus 2`."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_every_big_smem_kernel_on_both_devices():
    outs = []
    for dev in (0, 1, 0):
        torch.cuda.set_device(dev)
        g = torch.Generator(device=f"cuda:{dev}")
        g.manual_seed(11)
        x = torch.randn(2048, 4096, device=f"cuda:{dev}", generator=g)
        xb = x.to(torch.bfloat16)
        w = torch.rand(4096, device=f"cuda:{dev}", generator=g)
        res = [paddle.sum(x, axis=0),                       # TMA tensor-map column kernel (73 KB dynamic smem)
               F.softmax(x, axis=-1), F.log_softmax(xb, axis=-1),   # TMA row pipeline
               F.layer_norm(x, [4096], w, w),
               paddle.multiply_grad(x, w, x)[1],           # one-pass TMA multiply_grad
               paddle.sum(x), paddle.argmax(xb, axis=-1)]
        torch.cuda.synchronize(dev)
        outs.append([r.float().cpu().numpy() for r in res])
    for a, b, c in zip(*outs):
        assert np.array_equal(a, b) and np.array_equal(a, c)  # same seed, same kernels: identical on either device
