"""Two host threads, each on its own CUDA stream, interleave launches of DIFFERENT row sizes of the SAME kernel
instantiations (softmax / layer_norm row pipeline: the shared-memory size and carve-out are per-function attributes that
depend on the row size).  Every result must match the oracle: a launch must never run with the attributes another
thread has just set for its shape (csrc/common.cuh: FuncAttrHold)."""
import threading

import numpy as np
import pytest
import torch

from gpu_util import to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional


def test_two_threads_two_streams_interleaved_shapes(port):
    rng = np.random.default_rng(9)
    shapes = [(700, 2048), (600, 8192)]
    xs = [(rng.standard_normal(s) * 2).astype(np.float32) for s in shapes]
    ws = [rng.uniform(0.5, 1.5, s[1]).astype(np.float32) for s in shapes]
    bs = [rng.uniform(-0.5, 0.5, s[1]).astype(np.float32) for s in shapes]
    want_sm = [port.softmax(x, -1) for x in xs]
    want_ln = [port.layer_norm(x, 1, w, b, 1e-5)[0] for x, w, b in zip(xs, ws, bs)]
    dx = [to_dev(x) for x in xs]
    dw = [to_dev(w) for w in ws]
    db = [to_dev(b) for b in bs]
    torch.cuda.synchronize()
    errors = []
    start = threading.Barrier(2)

    def worker(which):
        try:
            torch.cuda.set_device(0)
            stream = torch.cuda.Stream()
            start.wait()
            with torch.cuda.stream(stream):
                for it in range(200):
                    k = (which + it) & 1 if it % 3 else which  # mostly alternate, sometimes repeat
                    y = F.softmax(dx[k], axis=-1)
                    z = F.layer_norm(dx[k], [shapes[k][1]], dw[k], db[k], 1e-5)
                    if it % 20 == 0:
                        stream.synchronize()
                        np.testing.assert_allclose(to_np(y), want_sm[k], rtol=1e-5, atol=1e-30)
                        np.testing.assert_allclose(to_np(z), want_ln[k], rtol=1e-5, atol=2e-5)
                stream.synchronize()
        except Exception as exc:  # noqa: BLE001
            errors.append(f"thread {which}: {exc!r}"[:600])

    ts = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    torch.cuda.synchronize()
    assert not errors, errors
