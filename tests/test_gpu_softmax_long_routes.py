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
