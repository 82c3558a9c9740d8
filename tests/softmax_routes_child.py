"""Child of tests/test_gpu_softmax_long_routes.py: the long-row softmax kernels, one forced route per process
(PDB200_SOFTMAX_LONG is read once per process).  Prints one JSON line {"ok": bool, "fails": [...]}."""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle_lib as O  # noqa: E402
from gpu_util import to_dev, to_np, ulp_diff_16  # noqa: E402

import paddlepaddle_b200 as paddle  # noqa: E402
from paddlepaddle_b200 import _lib as L  # noqa: E402

F = paddle.nn.functional
port = O.Oracle("port")
fails = []
rng = np.random.default_rng(int(os.environ.get("PDB200_SOFTMAX_LONG", "0")) + 100)
shapes = [(37, 40000), (5, 300000), (3, 1 << 20), (300, 16392), (2, 131072 + 8), (9, 128256)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in s.split("x")) for s in sys.argv[1:]]


def check32(tag, got, want, rtol=1e-5, atol=1e-30):
    if not np.allclose(got, want, rtol=rtol, atol=atol, equal_nan=True):
        bad = ~np.isclose(got, want, rtol=rtol, atol=atol, equal_nan=True)
        fails.append(f"{tag}: {int(bad.sum())} elements, first {np.argwhere(bad)[0].tolist()} got {got[bad][0]} want {want[bad][0]}")


def ref64(x):
    """float64 statement of the reference's formula (cpu/softmax_kernel.cc -> funcs/softmax_impl.h ValueClip: the shifted
    logits are clipped at -64): the restatement in oracle/ sums 10^6 float terms in index order, too coarse a yardstick
    for 4 MB rows."""
    x = x.astype(np.float64)
    with np.errstate(invalid="ignore"):
        d = np.maximum(x - x.max(axis=-1, keepdims=True), -64.0)
    e = np.exp(d)
    s = e.sum(axis=-1, keepdims=True)
    return e / s, d - np.log(s)


for shape in shapes:
    x = (rng.standard_normal(shape) * 3).astype(np.float32)
    x[0, :7] = -np.inf               # -inf elements: exp(-64) after the clip
    x[1, 5] = 200.0                  # one dominant element: everything else sits at the clip
    if shape[0] > 2:
        x[2, shape[1] // 2:] -= 80.0  # half the row below the clip relative to the other half
    w64, wl64 = ref64(x)
    want = w64.astype(np.float32)
    check32(f"softmax fp32 {shape}", to_np(F.softmax(to_dev(x), axis=-1)), want)
    check32(f"log_softmax fp32 {shape}", to_np(F.log_softmax(to_dev(x), axis=-1)), wl64.astype(np.float32), rtol=1e-5, atol=2e-5)
    for half, conv, back in (("bf16", port.f32_to_bf16, port.bf16_to_f32), ("fp16", port.f32_to_fp16, port.fp16_to_f32)):
        bits = conv(x)
        want_h = conv(ref64(back(bits))[0].astype(np.float32))
        got_h = to_np(F.softmax(to_dev(bits, half), axis=-1))
        d = ulp_diff_16(got_h, want_h)
        if d.max() > 1:
            fails.append(f"softmax {half} {shape}: {int((d > 1).sum())} elements beyond 1 ulp, max {int(d.max())}")
    # the forced route writes inside its output and nowhere else (compute-sanitizer is closed on this pool: canaries)
    for tdt in (torch.float32, torch.bfloat16):
        xt = to_dev(x).to(tdt)
        n_el, guard = xt.numel(), 4096
        buf = torch.empty(n_el + 2 * guard, dtype=tdt, device="cuda")
        buf.view(torch.uint8).fill_(0xA5)
        out = buf[guard:guard + n_el]
        lib = L.load()
        dd, rr = L.dims_array(shape)
        L.check(lib.pdb200_softmax(ctypes.c_void_p(xt.data_ptr()), dd, rr, -1, L.FLOAT32 if tdt == torch.float32 else L.BFLOAT16,
                                   ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), lib)
        torch.cuda.synchronize()
        raw, es = buf.view(torch.uint8), buf.element_size()
        if not (bool((raw[:guard * es] == 0xA5).all()) and bool((raw[(guard + n_el) * es:] == 0xA5).all())):
            fails.append(f"softmax {tdt} {shape}: wrote outside its output")
    # in place
    t = to_dev(x)
    F.softmax_(t, axis=-1)
    check32(f"softmax_ fp32 {shape}", to_np(t), want)
    # a row of -inf and a row with a NaN: NaN rows, the other rows untouched
    xs = x.copy()
    xs[0, :] = -np.inf
    xs[-1, shape[1] - 3] = np.nan
    got = to_np(F.softmax(to_dev(xs), axis=-1))
    if not (np.isnan(got[0]).all() and np.isnan(got[-1]).all()):
        fails.append(f"special rows {shape}: -inf row / NaN row not NaN")
    if shape[0] > 2:
        check32(f"special rows (others) {shape}", got[1:-1], want[1:-1])
torch.cuda.synchronize()
print(json.dumps({"ok": not fails, "fails": fails[:8]}))
