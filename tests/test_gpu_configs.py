"""BASELINE.json's configurations at FULL size on the GPU, checked against the CPU oracle (SURVEY.md §8d "Parity
sampling for big configs").

  C1  sum fp32 [4096,4096]                 every row vs the oracle
  C2  add bf16 / int32 [64,2048,4096]      the WHOLE tensor vs the oracle, slab by slab (bit-exact)
  C3  layer_norm bf16 [65536,8192]         EVERY row vs the oracle, both parameter dtypes
  C4  softmax fp32 / bf16 [32,32,2048,2048] >= 4096 seeded random rows + the first and last row of every 1/8 shard
  C5  sum / max of a 2^31-element shard    int64: every one of the 64 sub-shards exactly; fp32: every sub-shard against
                                           a float64 sum computed on the HOST from the data, 1e-5 * sum|x| of that
                                           sub-shard, plus the Eigen oracle on a 2^28 slice

The oracle is the compiled reference (oracle/_ref/libphi_cpu_ref.so — Paddle's own CPU kernels, shipped to the GPU box
prebuilt) when present, else the C restatement pinned to it.  The CPU kernels are single-threaded; rows / slabs /
sub-shards are independent, so the checks below fan disjoint slices out over host threads (ctypes releases the GIL) —
each call is still the reference's single-threaded kernel on its slice.  Integer sums and max are order-independent,
so adding up per-slice oracle results is exact.  Inputs come from seeded generators; nothing here reads /root/reference.
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

import oracle_lib as O
from gpu_util import assert_half_close, to_dev, to_np

pytestmark = pytest.mark.gpu

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional
THREADS = max(1, min(8, os.cpu_count() or 1))


@pytest.fixture(scope="module")
def cpu(port):
    """The strongest oracle available: the compiled reference if it was shipped, else the pinned restatement."""
    return O.Oracle("ref") if O.have_ref() else port


def _gen(shape, dtype, seed, scale=1.0):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    return (torch.randn(shape, device="cuda", dtype=torch.float32, generator=g) * scale).to(dtype)


def _pmap(fn, items):
    with ThreadPoolExecutor(THREADS) as ex:
        return list(ex.map(fn, items))


def test_c1_sum_fp32_4096(cpu):
    x = np.random.default_rng(1).random((4096, 4096), dtype=np.float32)
    got = to_np(paddle.sum(to_dev(x), axis=-1))
    np.testing.assert_allclose(got, cpu.reduce("sum", x, -1), rtol=1e-5)


def test_c2_add_bf16_whole_tensor(cpu):
    shape = (64, 2048, 4096)
    x = _gen(shape, torch.bfloat16, 2)
    y = _gen((4096,), torch.bfloat16, 22)
    out = paddle.add(x, y)
    assert out.shape == shape and out.dtype == torch.bfloat16
    yb = to_np(y)
    xv, ov = x.view(-1, 4096), out.view(-1, 4096)
    slab = 4096  # rows per oracle call: 32 MiB in, 32 MiB out

    def check(r0):
        want = cpu.elementwise("add", to_np(xv[r0:r0 + slab]), yb, O.BFLOAT16)
        return int((to_np(ov[r0:r0 + slab]) != want).sum())

    bad = sum(_pmap(check, range(0, xv.shape[0], slab)))
    assert bad == 0, f"config 2 bf16: {bad} of {xv.numel()} elements differ from the CPU kernel"


def test_c2_add_int32_whole_tensor(cpu):
    shape = (64, 2048, 4096)
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    xi = torch.randint(-2**30, 2**30, shape, device="cuda", dtype=torch.int32, generator=g)
    yi = torch.randint(-2**30, 2**30, (4096,), device="cuda", dtype=torch.int32, generator=g)
    oi = paddle.add(xi, yi)
    yb = to_np(yi)
    xv, ov = xi.view(-1, 4096), oi.view(-1, 4096)
    slab = 4096

    def check(r0):
        want = cpu.elementwise("add", to_np(xv[r0:r0 + slab]), yb)
        return int((to_np(ov[r0:r0 + slab]) != want).sum())

    bad = sum(_pmap(check, range(0, xv.shape[0], slab)))
    assert bad == 0, f"config 2 int32: {bad} elements differ from the CPU kernel"
    # and the exact integer checksum: sum(out) == sum(x) + rows * sum(y)   (mod 2^64)
    sx, so, sy = paddle.sum(xi), paddle.sum(oi), paddle.sum(yi)
    assert (int(so) - int(sx) - 131072 * int(sy)) % (1 << 64) == 0


def test_c3_layer_norm_bf16_every_row(cpu, port):
    rows, D = 65536, 8192
    r = torch.arange(rows, device="cuda").view(-1, 1)
    x = (_gen((rows, D), torch.float32, 4) * (1 + r % 7) + r % 3).to(torch.bfloat16)
    g = torch.Generator(device="cuda"); g.manual_seed(44)
    w = (torch.rand(D, device="cuda", generator=g) + 0.5)
    b = (torch.rand(D, device="cuda", generator=g) - 0.5)
    slab = 2048
    for wp, bp in ((w.to(torch.bfloat16), b.to(torch.bfloat16)), (w, b)):
        y, m, v = F.layer_norm(x, [D], wp, bp, 1e-5, return_stats=True)
        wf = port.bf16_to_f32(to_np(wp)) if wp.dtype == torch.bfloat16 else to_np(wp)
        bf = port.bf16_to_f32(to_np(bp)) if bp.dtype == torch.bfloat16 else to_np(bp)

        def check(r0):
            xs = port.bf16_to_f32(to_np(x[r0:r0 + slab]))
            oy, om, ov = cpu.layer_norm(xs, 1, wf, bf, 1e-5)
            # 1 bf16 ulp; atol: outputs near zero are small differences of O(1) terms (see test_gpu_parity.py::test_layer_norm_routes)
            assert_half_close(to_np(y[r0:r0 + slab]), port.f32_to_bf16(oy), 1, f"config 3 rows {r0}..", atol=2e-5)
            np.testing.assert_allclose(to_np(m[r0:r0 + slab]), om, rtol=1e-5, atol=1e-5 * float(np.abs(xs).max()))
            # variance: 1e-5 against the float64 value; the CPU kernel itself adds 8192 squares one after the other in
            # float32 (RowwiseMean2D, funcs/math_function_impl.h:158-191) and sits up to ~4e-5 from it on some rows
            v64 = x[r0:r0 + slab].double().var(dim=1, unbiased=False).cpu().numpy()
            np.testing.assert_allclose(to_np(v[r0:r0 + slab]), v64, rtol=1e-5)
            np.testing.assert_allclose(to_np(v[r0:r0 + slab]), ov, rtol=1e-4)
            return slab

        assert sum(_pmap(check, range(0, rows, slab))) == rows


def _c4_rows(n_rows, shards=8, n_random=4096, seed=55):
    per = n_rows // shards
    edge = [s * per for s in range(shards)] + [(s + 1) * per - 1 for s in range(shards)]
    rnd = np.random.default_rng(seed).choice(n_rows, n_random, replace=False)
    return np.unique(np.concatenate([np.asarray(edge, np.int64), rnd.astype(np.int64)]))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_c4_softmax_full(cpu, port, dtype):
    shape = (32, 32, 2048, 2048)
    x = _gen(shape, dtype, 5, scale=2.0)
    y = F.softmax(x, axis=-1)
    flat_x, flat_y = x.view(-1, 2048), y.view(-1, 2048)
    rows = _c4_rows(flat_x.shape[0])
    assert len(rows) >= 4096
    idx = torch.from_numpy(rows).cuda()
    xs, ys = to_np(flat_x[idx]), to_np(flat_y[idx])
    if dtype == torch.float32:
        np.testing.assert_allclose(ys, cpu.softmax(xs, -1), rtol=1e-5)
    else:
        want = port.f32_to_bf16(cpu.softmax(port.bf16_to_f32(xs), -1))
        assert_half_close(ys, want, 1, "config 4 bf16 sampled rows")
    del x
    # property over every row: sums to 1 (fp32: to 1e-5; bf16 outputs are each rounded to 2^-9 relative)
    s = paddle.sum(flat_y, axis=-1, dtype="float32") if dtype != torch.float32 else paddle.sum(flat_y, axis=-1)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert float((s.float() - 1).abs().max()) < tol


def _sub_shards(x, n_sub):
    """Yield (i, host numpy array) for the n_sub equal sub-shards of the flat device tensor x, through two pinned
    staging buffers (the shard itself never sits on the host whole)."""
    per = x.numel() // n_sub
    bufs = [torch.empty(per, dtype=x.dtype, pin_memory=True) for _ in range(2)]
    for i in range(n_sub):
        b = bufs[i & 1]
        b.copy_(x[i * per:(i + 1) * per])
        yield i, b.numpy()


def test_c5_int64_every_sub_shard_exact(cpu):
    """One rank's shard of config 5: 2^31 int64 (16 GiB).  Every one of 64 sub-shards is summed / maxed by the CPU
    oracle (8 host threads over disjoint slices of a sub-shard: integer sum and max are order-independent) and must
    equal the GPU's row results exactly; the wrapped total and the overall max and argmax follow from those."""
    n, n_sub = 1 << 31, 64
    g = torch.Generator(device="cuda"); g.manual_seed(6)
    x = torch.randint(-2**40, 2**40, (n,), device="cuda", dtype=torch.int64, generator=g)
    total, mx, am = paddle.sum(x), paddle.max(x), paddle.argmax(x)
    parts = to_np(paddle.sum(x.view(n_sub, -1), axis=1))
    pmax = to_np(paddle.max(x.view(n_sub, -1), axis=1))
    host_total, host_max, host_argmax = 0, None, None
    per = n // n_sub
    for i, xs in _sub_shards(x, n_sub):
        pieces = np.array_split(xs, THREADS)
        sums = _pmap(lambda p: int(cpu.reduce("sum", p)), pieces)
        maxs = _pmap(lambda p: int(cpu.reduce("max", p)), pieces)
        s = sum(sums) % (1 << 64)
        s = s - (1 << 64) if s >= (1 << 63) else s
        assert int(parts[i]) == s, f"sub-shard {i}: sum differs"
        assert int(pmax[i]) == max(maxs), f"sub-shard {i}: max differs"
        host_total = (host_total + s) % (1 << 64)
        if host_max is None or max(maxs) > host_max:
            host_max = max(maxs)
            host_argmax = i * per + int(np.argmax(xs))  # first occurrence inside the first sub-shard holding the max
    host_total = host_total - (1 << 64) if host_total >= (1 << 63) else host_total
    assert int(total) == host_total      # exact, wraps identically
    assert int(mx) == host_max
    assert int(am) == host_argmax        # lowest index among ties


def test_c5_fp32_every_sub_shard(cpu):
    """2^31 fp32 (8 GiB).  Reference = float64 sums computed on the host from the data itself; tolerance 1e-5 * sum|x| of
    the sub-shard (north_star: 1e-5 relative for fp32 reductions accumulated in fp32; sums with cancellation are judged
    against sum|x|).  Max is exact.  The first 2^28 elements are also summed by the Eigen oracle (fp32, one thread)."""
    n, n_sub = 1 << 31, 64
    g = torch.Generator(device="cuda"); g.manual_seed(6)
    x = torch.rand((n,), device="cuda", dtype=torch.float32, generator=g) * 2 - 1
    total, mx, am = paddle.sum(x), paddle.max(x), paddle.argmax(x)
    parts = to_np(paddle.sum(x.view(n_sub, -1), axis=1))
    pmax = to_np(paddle.max(x.view(n_sub, -1), axis=1))
    ref_total, abs_total, host_max, host_argmax = 0.0, 0.0, None, None
    per = n // n_sub
    for i, xs in _sub_shards(x, n_sub):
        pieces = np.array_split(xs, THREADS)
        s64 = sum(_pmap(lambda p: float(np.sum(p, dtype=np.float64)), pieces))
        a64 = sum(_pmap(lambda p: float(np.sum(np.abs(p), dtype=np.float64)), pieces))
        assert abs(float(parts[i]) - s64) <= 1e-5 * a64, f"sub-shard {i}: {parts[i]} vs {s64}"
        m = float(xs.max())
        assert float(pmax[i]) == m, f"sub-shard {i}: max differs"
        ref_total += s64
        abs_total += a64
        if host_max is None or m > host_max:
            host_max, host_argmax = m, i * per + int(np.argmax(xs))
    assert abs(float(total) - ref_total) <= 1e-5 * abs_total
    assert float(mx) == host_max
    assert int(am) == host_argmax
    # the Eigen oracle (fp32 accumulation, its own order) on a 2^28 slice
    sl = x[:1 << 28]
    want = float(cpu.reduce("sum", to_np(sl)))
    got = float(paddle.sum(sl))
    abs_sl = float(sl.abs().double().sum())
    # both sides are fp32 accumulations of 2^28 terms: each is within its own rounding of the true sum
    assert abs(got - want) <= 1e-5 * abs_sl
