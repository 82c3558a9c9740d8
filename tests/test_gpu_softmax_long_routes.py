"""Every long-row softmax kernel (csrc/softmax.cu) forced in its own process — PDB200_SOFTMAX_LONG is read once per
process — against the restatement: 1 the cluster route (lean shared-memory kernel), 2 two sweeps with the second out of L2,
5 the first cluster kernel (online rescale), 6 the relay kernel (persistent CTAs, chunk records in global memory), 0 the
default choice.  Shapes cover rows of 80 KB to 4 MB, a ragged last chunk, rows of -inf, NaN, elements below the -64 clip, in
place, softmax and log_softmax, float32 / bfloat16 / float16."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("route", [0, 1, 2, 5, 6])
def test_long_row_routes(route):
    env = dict(os.environ, PDB200_SOFTMAX_LONG=str(route))
    r = subprocess.run([sys.executable, os.path.join(HERE, "softmax_routes_child.py")], capture_output=True, text=True,
                       timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["ok"], out["fails"]


@pytest.mark.parametrize("cfg", [1, 3, 6])
def test_relay_geometries(cfg):
    """The relay kernel's CTA shapes (256 x 32, 512 x 32, 128 x 32 elements per chunk), each forced on all shapes."""
    env = dict(os.environ, PDB200_SOFTMAX_LONG="6", PDB200_RELAY_CFG=str(cfg))
    r = subprocess.run([sys.executable, os.path.join(HERE, "softmax_routes_child.py"), "37x40000", "3x1048576", "9x128256", "300x16392"],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["ok"], out["fails"]


def test_relay_inside_a_cuda_graph():
    """The default long-row route for float32 (cooperative launch, stream-ordered record array) captured into a CUDA
    graph and replayed on new data: same bits as the eager call."""
    import torch

    import paddlepaddle_b200 as paddle
    F = paddle.nn.functional
    x = torch.randn(1024, 40000, device="cuda")  # 164 MB of float32: the default route is the relay kernel
    y = torch.empty_like(x)
    eager = F.softmax(x, axis=-1)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        F.softmax(x, axis=-1)  # warm-up on the capture stream
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        y.copy_(F.softmax(x, axis=-1))
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(y, eager)
    x.copy_(torch.randn(1024, 40000, device="cuda") * 3)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(y, F.softmax(x, axis=-1))


@pytest.mark.parametrize("shape,dt", [((8, 2048, 4096), "bfloat16"), ((24, 1000, 2056), "float32"), ((40, 300, 4104), "float16")])
def test_strided_axis_cluster_route(shape, dt):
    """softmax / log_softmax along a middle axis of a tensor larger than L2: the column tile is held in the shared memory
    of a cluster (csrc/softmax.cu softmax_cols_cluster_kernel) and x is read once.  Against a float64 statement of the
    reference's formula (shifted logits clipped at -64); ragged row slices (1000 = 4 x 250, 300 = 2 x 150), a partial last
    column tile, a column of -inf, a NaN, columns far below the clip."""
    import numpy as np
    import torch

    import paddlepaddle_b200 as paddle
    from gpu_util import ulp_diff_16
    F = paddle.nn.functional
    tdt = getattr(torch, dt)
    g = torch.Generator(device="cuda")
    g.manual_seed(7)
    x = (torch.randn(shape, device="cuda", generator=g) * 3).to(tdt)
    x[0, :, 5] = float("-inf")          # a whole column of -inf -> NaN column
    x[1, 7, 9] = float("nan")           # one NaN -> NaN column
    x[2, : shape[1] // 2, 11] -= 100.0  # half a column far below the clip
    x[3, 3, :] = 50.0                   # one dominant row
    y = F.softmax(x, axis=1)
    ly = F.log_softmax(x, axis=1)
    for o in (0, 1, 2, 3, shape[0] - 1):  # whole [dim, inner] slabs against float64
        xs = x[o].double().cpu().numpy()
        with np.errstate(invalid="ignore"):
            d = np.maximum(xs - xs.max(axis=0, keepdims=True), -64.0)
        e = np.exp(d)
        s = e.sum(axis=0, keepdims=True)
        want, lwant = e / s, d - np.log(s)
        got, lgot = y[o].float().cpu().numpy(), ly[o].float().cpu().numpy()
        nanc = np.isnan(want).any(axis=0)
        assert np.isnan(got[:, nanc]).all() and np.isnan(lgot[:, nanc]).all()
        if dt == "float32":
            np.testing.assert_allclose(got[:, ~nanc], want[:, ~nanc], rtol=1e-5, atol=1e-30)
            np.testing.assert_allclose(lgot[:, ~nanc], lwant[:, ~nanc], rtol=1e-5, atol=2e-5)
        else:
            wt = torch.from_numpy(want[:, ~nanc]).to(tdt).view(torch.int16).numpy().view(np.uint16)
            gt = y[o][:, torch.from_numpy(~nanc).cuda()].contiguous().view(torch.int16).cpu().numpy().view(np.uint16)
            assert ulp_diff_16(gt, wt).max() <= 1
            np.testing.assert_allclose(lgot[:, ~nanc], lwant[:, ~nanc], rtol=1e-2, atol=5e-2)
    # in place
    x2 = x.clone()
    F.softmax_(x2, axis=1)
    assert torch.equal(torch.nan_to_num(x2.float(), nan=-1.0), torch.nan_to_num(y.float(), nan=-1.0))
