"""CPU-only checks: the C-ABI library loads, exports every symbol include/pdb200.h declares, and the
host-side planning / argument validation behaves like the reference's InferMeta + PADDLE_ENFORCE
checks.  No compute call needs a GPU here: validation happens before any CUDA call."""
import ctypes
import os
import re

import numpy as np
import pytest

from paddlepaddle_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    return L.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "pdb200.h")).read()
    declared = set(re.findall(r"PDB200_API[^;(]*?\b(pdb200_\w+)\s*\(", hdr))
    assert len(declared) >= 16
    raw = ctypes.CDLL(L.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in include/pdb200.h but not exported"
    assert declared == set(L.SIGNATURES), "ctypes table and header disagree"
    assert b"sm_100a" in lib.pdb200_version()


def _bshape(lib, xs, ys):
    xd, xr = L.dims_array(xs)
    yd, yr = L.dims_array(ys)
    od = (ctypes.c_int64 * 9)()
    r = ctypes.c_int(0)
    rc = lib.pdb200_broadcast_shape(xd, xr, yd, yr, od, ctypes.byref(r))
    return rc, tuple(od[i] for i in range(r.value))


@pytest.mark.parametrize("xs,ys", [((2, 10, 12), (12,)), ((10, 3, 1, 4), (10, 1, 12, 1)), ((10, 12), (2, 2, 10, 12)),
                                   ((), (3,)), ((), ()), ((5, 0, 3), (1, 3)), ((1, 3), (5, 0, 3)), ((64, 2048, 4096), (4096,))])
def test_broadcast_shape_matches_numpy(lib, xs, ys):
    rc, out = _bshape(lib, xs, ys)
    assert rc == L.OK and out == np.broadcast_shapes(xs, ys)


def test_broadcast_mismatch_is_invalid_argument(lib):
    rc, _ = _bshape(lib, (2, 3), (4,))
    assert rc == L.INVALID_ARGUMENT
    msg = lib.pdb200_last_error().decode()
    assert "Broadcast dimension mismatch" in msg and "[2, 3]" in msg
    with pytest.raises(ValueError):
        L.check(rc, lib)


def test_elementwise_rejects_bad_shapes_before_touching_cuda(lib):
    xd, xr = L.dims_array((2, 3))
    yd, yr = L.dims_array((4,))
    rc = lib.pdb200_add(None, xd, xr, None, yd, yr, None, L.FLOAT32, None)
    assert rc == L.INVALID_ARGUMENT
    # empty output: nothing to do, no pointers needed
    xd, xr = L.dims_array((0, 3))
    yd, yr = L.dims_array((3,))
    assert lib.pdb200_add(None, xd, xr, None, yd, yr, None, L.FLOAT32, None) == L.OK


def test_reduce_validation(lib):
    xd, xr = L.dims_array((4, 5))
    ad, an = L.dims_array((2,))
    rc = lib.pdb200_sum_raw(None, xd, xr, ad, an, 0, L.FLOAT32, L.FLOAT32, None, None, 0, None)
    assert rc == L.INVALID_ARGUMENT and "range" in lib.pdb200_last_error().decode()
    xd0, xr0 = L.dims_array((4, 0))
    ad, an = L.dims_array((1,))
    rc = lib.pdb200_max_raw(None, xd0, xr0, ad, an, 0, L.FLOAT32, L.FLOAT32, None, None, 0, None)
    assert rc == L.INVALID_ARGUMENT and "must not empty" in lib.pdb200_last_error().decode()
    rc = lib.pdb200_prod(None, xd, xr, ad, an, 0, L.BOOL, L.BOOL, None, None, 0, None)
    assert rc == L.UNSUPPORTED


def test_workspace_queries(lib):
    def ws(op, shape, axes, xdt=L.FLOAT32, odt=L.FLOAT32):
        xd, xr = L.dims_array(shape)
        ad, an = L.dims_array(axes)
        return lib.pdb200_reduce_workspace_bytes(op, xd, xr, ad, an, int(not axes), xdt, odt)
    assert ws(L.REDUCE_SUM, (4096, 4096), (-1,)) == 0          # config 1: one pass, no scratch
    full = ws(L.REDUCE_SUM, (1 << 26,), ())                     # full reduce: partials
    assert 0 < full <= 1 << 20
    assert ws(L.REDUCE_SUM, (8, 16, 32, 64), (0, 2)) > 0        # two reduce runs -> intermediate
    assert ws(L.REDUCE_MAX, (1 << 26,), (), L.INT64, L.INT64) > 0
    xd, xr = L.dims_array((1 << 26,))
    assert lib.pdb200_argmax_workspace_bytes(xd, xr, 0, 1, L.FLOAT32) > 0
    xd, xr = L.dims_array((1024, 64))
    assert lib.pdb200_argmax_workspace_bytes(xd, xr, 1, 0, L.FLOAT32) == 0


def test_softmax_layernorm_validation(lib):
    d, r = L.dims_array((4, 5))
    assert lib.pdb200_softmax(None, d, r, 2, L.FLOAT32, None, None) == L.INVALID_ARGUMENT
    assert "axis" in lib.pdb200_last_error().decode()
    d0, r0 = L.dims_array((0, 5))
    assert lib.pdb200_softmax(None, d0, r0, -1, L.FLOAT32, None, None) == L.OK  # 0-size: no-op
    assert lib.pdb200_softmax(None, d, r, -1, L.INT32, None, None) in (L.UNSUPPORTED, L.INVALID_ARGUMENT)
    rc = lib.pdb200_layer_norm(None, d, r, None, None, L.FLOAT32, 1e-5, 3, L.FLOAT32, None, None, None, None)
    assert rc == L.INVALID_ARGUMENT and "begin_norm_axis" in lib.pdb200_last_error().decode()


def test_python_api_refuses_cpu_tensors():
    import torch

    import paddlepaddle_b200 as paddle
    with pytest.raises(RuntimeError, match="no CPU kernels"):
        paddle.sum(torch.ones(4))
    with pytest.raises(RuntimeError, match="no CPU kernels"):
        paddle.nn.functional.softmax(torch.ones(4, 4))


def test_product_never_touches_the_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "paddlepaddle_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in src.lower().replace("oracle is built from", "").replace("parity oracle", ""), f
    out = os.popen(f"ldd {L.LIB_PATH}").read()
    assert "oracle" not in out and "phi_cpu_ref" not in out
