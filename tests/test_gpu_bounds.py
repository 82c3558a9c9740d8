"""Out-of-bounds guard: compute-sanitizer is closed on this GPU pool, so every output of every route is
written into the middle of a canary-filled buffer and the canaries are checked afterwards.  The C-ABI is
called directly (ctypes) so that the output pointer is ours, not an allocator's."""
import ctypes

import numpy as np
import pytest
import torch

from paddlepaddle_b200 import _lib as L

pytestmark = pytest.mark.gpu

GUARD = 4096  # elements of canary on each side
CODES = {torch.float32: L.FLOAT32, torch.bfloat16: L.BFLOAT16, torch.int64: L.INT64, torch.int32: L.INT32,
         torch.float16: L.FLOAT16, torch.float64: L.FLOAT64}


def _guarded(numel, dtype):
    """(whole buffer, view of `numel` elements in the middle, 16-byte aligned)."""
    buf = torch.empty(numel + 2 * GUARD, dtype=dtype, device="cuda")
    buf.view(torch.uint8).fill_(0xA5)
    return buf, buf[GUARD:GUARD + numel]


def _check(buf, numel, what):
    raw = buf.view(torch.uint8)
    es = buf.element_size()
    assert bool((raw[:GUARD * es] == 0xA5).all()), f"{what}: wrote BEFORE the output"
    assert bool((raw[(GUARD + numel) * es:] == 0xA5).all()), f"{what}: wrote PAST the output"


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _st():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("xs,ys", [((257, 4096), (4096,)), ((300, 1000), (1000,)), ((33, 1001), (1001,)), ((512, 2048), (512, 1)),
                                   ((1000003,), (1000003,)), ((2, 3, 4, 5, 6, 7), (3, 1, 5, 1, 7)), ((1, 100000), (1,)), ((7, 130), (7, 1))])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64])
def test_elementwise_stays_in_bounds(xs, ys, dtype):
    lib = L.load()
    x = torch.ones(xs, device="cuda").to(dtype)
    y = torch.ones(ys, device="cuda").to(dtype)
    oshape = np.broadcast_shapes(xs, ys)
    n = int(np.prod(oshape))
    buf, out = _guarded(n, dtype)
    xd, xr = L.dims_array(xs)
    yd, yr = L.dims_array(ys)
    for fn in (lib.pdb200_add, lib.pdb200_subtract):
        L.check(fn(_p(x), xd, xr, _p(y), yd, yr, _p(out), CODES[dtype], _st()), lib)
        torch.cuda.synchronize()
        _check(buf, n, f"elementwise {xs} {ys} {dtype}")
    assert float(out.float().abs().max()) == 0.0  # subtract of ones: every element written


@pytest.mark.parametrize("shape,axes", [((4096, 4096), (1,)), ((1000, 3), (1,)), ((37, 5000), (1,)), ((3, 70000), (1,)),
                                        ((4096, 512), (0,)), ((70000, 24), (0,)), ((129, 7, 33), (1,)), ((8, 16, 32, 64), (0, 2)),
                                        ((1 << 22,), ())])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64])
def test_reduce_and_argmax_stay_in_bounds(shape, axes, dtype):
    lib = L.load()
    x = torch.ones(shape, device="cuda").to(dtype)
    red = set(axes) if axes else set(range(len(shape)))
    n = int(np.prod([d for i, d in enumerate(shape) if i not in red])) if len(red) < len(shape) else 1
    xd, xr = L.dims_array(shape)
    ad, an = L.dims_array(axes)
    code = CODES[dtype]
    ra = int(not axes)
    for op, fn in ((L.REDUCE_SUM, lib.pdb200_sum_raw), (L.REDUCE_MAX, lib.pdb200_max_raw)):
        nbytes = lib.pdb200_reduce_workspace_bytes(op, xd, xr, ad, an, ra, code, code)
        wbuf, ws = _guarded(max(nbytes, 16), torch.uint8)
        buf, out = _guarded(n, dtype)
        L.check(fn(_p(x), xd, xr, ad, an, ra, code, code, _p(out), _p(ws), nbytes, _st()), lib)
        torch.cuda.synchronize()
        _check(buf, n, f"reduce out {shape} {axes} {dtype}")
        _check(wbuf, max(nbytes, 16), f"reduce workspace {shape} {axes} {dtype}")
    if len(axes) <= 1:
        flatten = int(not axes)
        ax = axes[0] if axes else 0
        nbytes = lib.pdb200_argmax_workspace_bytes(xd, xr, ax, flatten, code)
        wbuf, ws = _guarded(max(nbytes, 16), torch.uint8)
        buf, out = _guarded(n, torch.int64)
        L.check(lib.pdb200_argmax(_p(x), xd, xr, ax, flatten, code, L.INT64, _p(out), _p(ws), nbytes, _st()), lib)
        torch.cuda.synchronize()
        _check(buf, n, f"argmax out {shape} {axes} {dtype}")
        _check(wbuf, max(nbytes, 16), f"argmax workspace {shape} {axes}")
        assert int(out.max()) == 0  # all ones: the first index everywhere


@pytest.mark.parametrize("rows,D", [(600, 2048), (33, 1001), (1000, 24), (7, 4096), (600, 8192), (3, 16384), (2, 40000),
                                    (40, 131072), (300, 20000), (2, 300000)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_softmax_and_layer_norm_stay_in_bounds(rows, D, dtype):
    lib = L.load()
    x = torch.randn((rows, D), device="cuda").to(dtype)
    w = torch.ones(D, device="cuda").to(dtype)
    b = torch.zeros(D, device="cuda").to(dtype)
    d, r = L.dims_array((rows, D))
    code = CODES[dtype]
    buf, out = _guarded(rows * D, dtype)
    L.check(lib.pdb200_softmax(_p(x), d, r, -1, code, _p(out), _st()), lib)
    torch.cuda.synchronize()
    _check(buf, rows * D, f"softmax {rows}x{D} {dtype}")
    buf, out = _guarded(rows * D, dtype)
    mbuf, m = _guarded(rows, torch.float32)
    vbuf, v = _guarded(rows, torch.float32)
    L.check(lib.pdb200_layer_norm(_p(x), d, r, _p(w), _p(b), code, ctypes.c_float(1e-5), 1, code, _p(out), _p(m), _p(v), _st()), lib)
    torch.cuda.synchronize()
    _check(buf, rows * D, f"layer_norm y {rows}x{D} {dtype}")
    _check(mbuf, rows, "layer_norm mean")
    _check(vbuf, rows, "layer_norm variance")
    # non-last-axis softmax
    x3 = torch.randn((5, 37, 129), device="cuda").to(dtype)
    d3, r3 = L.dims_array((5, 37, 129))
    buf, out = _guarded(x3.numel(), dtype)
    L.check(lib.pdb200_softmax(_p(x3), d3, r3, 1, code, _p(out), _st()), lib)
    torch.cuda.synchronize()
    _check(buf, x3.numel(), "softmax axis=1")


@pytest.mark.parametrize("shape,axes", [((4096, 512), (1,)), ((4096, 512), (0,)), ((257, 1031), (0,)), ((37, 96, 40), (1,)),
                                        ((16, 13, 2048), (1,)), ((7, 33, 5, 16), (1, 3)), ((3, 70000), (1,)), ((1000003,), (0,))])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64])
def test_reduction_backward_stays_in_bounds(shape, axes, dtype):
    """max_grad / prod_grad / sum_grad on every route (outer, inner, batched inner, general; ragged and unaligned)."""
    lib = L.load()
    x = torch.ones(shape, device="cuda").to(dtype)
    oshape = [s for i, s in enumerate(shape) if i not in axes]
    out = torch.ones(oshape, device="cuda").to(dtype)
    og = torch.full(oshape, 3, device="cuda").to(dtype)
    n = x.numel()
    xd, xr = L.dims_array(shape)
    ad, an = L.dims_array(axes)
    code = CODES[dtype]
    for name in ("pdb200_max_grad", "pdb200_prod_grad"):
        buf, dx = _guarded(n, dtype)
        L.check(getattr(lib, name)(_p(x), xd, xr, _p(out), _p(og), ad, an, 0, code, _p(dx), _st()), lib)
        torch.cuda.synchronize()
        _check(buf, n, f"{name} {shape} {axes} {dtype}")
        assert bool((dx == 3).all()), f"{name}: every element written (x == out == 1 everywhere)"
    buf, dx = _guarded(n, dtype)
    L.check(lib.pdb200_sum_grad(_p(og), xd, xr, ad, an, 0, code, _p(dx), _st()), lib)
    torch.cuda.synchronize()
    _check(buf, n, f"sum_grad {shape} {axes} {dtype}")
    assert bool((dx == 3).all())


@pytest.mark.parametrize("rows,D", [(700, 1024), (33, 1000), (7, 100), (600, 8192), (5, 20000)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fused_layer_norm_stays_in_bounds(rows, D, dtype):
    lib = L.load()
    x = torch.randn((rows, D), device="cuda").to(dtype)
    res = torch.randn((rows, D), device="cuda").to(dtype)
    bias = torch.randn(D, device="cuda").to(dtype)
    w = torch.ones(D, device="cuda")
    b = torch.zeros(D, device="cuda")
    d, r = L.dims_array((rows, D))
    code = CODES[dtype]
    obuf, out = _guarded(rows * D, dtype)
    rbuf, ro = _guarded(rows * D, dtype)
    mbuf, m = _guarded(rows, torch.float32)
    vbuf, v = _guarded(rows, torch.float32)
    L.check(lib.pdb200_fused_bias_residual_layernorm(_p(x), d, r, _p(bias), _p(res), _p(w), _p(b), ctypes.c_float(1e-5),
                                                     ctypes.c_float(0.5), 1, code, _p(out), _p(ro), _p(m), _p(v), _st()), lib)
    torch.cuda.synchronize()
    for bb, nn, what in ((obuf, rows * D, "out"), (rbuf, rows * D, "residual_out"), (mbuf, rows, "mean"), (vbuf, rows, "variance")):
        _check(bb, nn, f"fused_layer_norm {what} {rows}x{D} {dtype}")
    obuf, out = _guarded(rows * D, dtype)  # the sum-only form
    L.check(lib.pdb200_fused_bias_residual_layernorm(_p(x), d, r, _p(bias), _p(res), None, None, ctypes.c_float(1e-5),
                                                     ctypes.c_float(0.5), 1, code, _p(out), None, None, None, _st()), lib)
    torch.cuda.synchronize()
    _check(obuf, rows * D, f"fused_layer_norm sum-only {rows}x{D} {dtype}")
