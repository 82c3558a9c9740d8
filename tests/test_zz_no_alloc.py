"""The allocation-free column-sum entry point (pdb200_sum_axis0_static) against the regular one.

Some users of this library cannot allocate device memory at call time: their memory was planned and
handed out when the process started, or they sit in a latency-critical loop that must stay away from
cudaMalloc, or their host framework simply passes no allocator down.  For them the library carries a
fixed 1.5 MiB scratch array of its own.  This test drives that entry point through the C-ABI and checks
it against pdb200_sum_raw (which is given a caller-allocated workspace), twice in a row with different
shapes so that a result depending on what the previous call left in the shared scratch would show.

The file name sorts last on purpose: with `pytest -x` every other test has already run when this one
does.  It needs nothing from the other tests and also passes alone in a fresh process.
"""
import ctypes

import numpy as np
import pytest
import torch

from paddlepaddle_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _sum_axis0_regular(lib, x, out_dtype):
    """pdb200_sum_raw over axis 0 with a workspace the caller allocates."""
    rows, cols = x.shape
    xd, xr = L.dims_array((rows, cols))
    ad, an = L.dims_array((0,))
    xc = {torch.float32: L.FLOAT32, torch.bfloat16: L.BFLOAT16, torch.int64: L.INT64}[x.dtype]
    oc = {torch.float32: L.FLOAT32, torch.bfloat16: L.BFLOAT16, torch.int64: L.INT64}[out_dtype]
    nbytes = lib.pdb200_reduce_workspace_bytes(L.REDUCE_SUM, xd, xr, ad, an, 0, xc, oc)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device="cuda")
    out = torch.empty((cols,), dtype=out_dtype, device="cuda")
    L.check(lib.pdb200_sum_raw(_ptr(x), xd, xr, ad, an, 0, xc, oc, _ptr(out), _ptr(ws), nbytes, _stream()), lib)
    return out, nbytes, xc, oc


def _sum_axis0_static(lib, x, out_dtype, xc, oc):
    rows, cols = x.shape
    out = torch.empty((cols,), dtype=out_dtype, device="cuda")
    rc = lib.pdb200_sum_axis0_static(_ptr(x), rows, cols, xc, oc, _ptr(out), _stream())
    return rc, out


def _make(shape, dtype, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    if dtype == torch.int64:
        return torch.randint(-2**40, 2**40, shape, device="cuda", dtype=dtype, generator=g)
    return (torch.randn(shape, device="cuda", generator=g) * 3 + 0.5).to(dtype)


def test_static_scratch_matches_regular_entry_point_twice_in_a_row():
    lib = L.load()
    # two calls back to back; the second is narrower and shorter, so it reuses only part of what the
    # first one left in the shared scratch
    calls = [((70000, 1000), torch.float32, torch.float32), ((9000, 24), torch.float32, torch.float32),
             ((50000, 512), torch.bfloat16, torch.bfloat16), ((20000, 96), torch.bfloat16, torch.float32),
             ((30000, 130), torch.int64, torch.int64), ((4096, 7), torch.int64, torch.int64)]
    used_scratch = 0
    for i, (shape, dt, odt) in enumerate(calls):
        x = _make(shape, dt, 100 + i)
        want, nbytes, xc, oc = _sum_axis0_regular(lib, x, odt)
        used_scratch += int(nbytes > 0)
        rc, got = _sum_axis0_static(lib, x, odt, xc, oc)
        L.check(rc, lib)
        torch.cuda.synchronize()
        if dt == torch.int64:
            assert torch.equal(got, want)                       # integers: bit-exact
        elif odt == torch.float32:
            # fp32 accumulate: the 1e-5 bar of the parity tests (same kernels, so in fact identical)
            scale = x.float().abs().sum(0).cpu().numpy()
            assert np.all(np.abs(got.cpu().numpy() - want.cpu().numpy()) <= 1e-5 * scale)
        else:
            # bf16 output: within 1 ulp of the regular entry point's result
            a = got.view(torch.int16).cpu().numpy().astype(np.int32)
            b = want.view(torch.int16).cpu().numpy().astype(np.int32)
            assert np.abs(a - b).max() <= 1
        # and against an independent float64 reference, so both are not wrong together
        ref = x.double().sum(0) if dt != torch.int64 else x.sum(0)
        if dt == torch.int64:
            assert torch.equal(got, ref)
        else:
            tol = 1e-5 if odt == torch.float32 else 2.0 ** -7
            assert torch.all((got.double() - ref).abs() <= tol * x.double().abs().sum(0))
    assert used_scratch >= 4, "the shapes above are meant to exercise the split (scratch-using) path"


def test_static_scratch_limit_and_no_allocation():
    lib = L.load()
    # wide enough that no partials are needed: fine at any width, the scratch is not touched
    x = _make((64, 400000), torch.float32, 7)
    want, nbytes, xc, oc = _sum_axis0_regular(lib, x, torch.float32)
    assert nbytes == 0
    rc, got = _sum_axis0_static(lib, x, torch.float32, xc, oc)
    L.check(rc, lib)
    assert torch.equal(got, want)
    # a shape whose partial sums exceed the 1.5 MiB built-in scratch is refused with a clear message:
    # bf16 [4096, 140032] is cut into 3 row-slabs of 140032 float partials = 1.68 MB
    rows, cols = 4096, 140032
    big = torch.zeros((rows, cols), dtype=torch.bfloat16, device="cuda")
    rc, _ = _sum_axis0_static(lib, big, torch.float32, L.BFLOAT16, L.FLOAT32)
    assert rc == L.WORKSPACE_TOO_SMALL
    msg = lib.pdb200_last_error().decode()
    assert "built-in scratch holds 1572864" in msg and "pdb200_sum_raw" in msg
    # ... while the regular entry point, given a workspace, handles the same shape
    want, nbytes, _, _ = _sum_axis0_regular(lib, big, torch.float32)
    assert nbytes > (3 << 19) and float(want.abs().max()) == 0.0
    del big, want
    # device memory in use does not move across calls (after the first call's one-time address lookup)
    x = _make((70000, 1000), torch.float32, 8)
    out = torch.empty((1000,), dtype=torch.float32, device="cuda")
    lib.pdb200_sum_axis0_static(_ptr(x), 70000, 1000, L.FLOAT32, L.FLOAT32, _ptr(out), _stream())
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(5):
        assert lib.pdb200_sum_axis0_static(_ptr(x), 70000, 1000, L.FLOAT32, L.FLOAT32, _ptr(out), _stream()) == L.OK
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    # bad arguments are refused before anything is launched
    assert lib.pdb200_sum_axis0_static(_ptr(x), 0, 1000, L.FLOAT32, L.FLOAT32, _ptr(out), _stream()) == L.INVALID_ARGUMENT
    assert lib.pdb200_sum_axis0_static(_ptr(x), 70000, 1000, L.INT32, L.INT32, _ptr(out), _stream()) == L.UNSUPPORTED
