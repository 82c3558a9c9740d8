"""The NVLink peer-memory exchange (csrc/peer.cu, PeerGroup) on two GPUs, one process each: results must equal the NCCL
AllReduce and the gathered partials, over many calls (the two slots alternate), for scalars and vectors, every op and
dtype.  Skipped on a single-GPU box; run with `gpurun --gpus 2`."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                          LOCAL_RANK=str(rank), PDB200_PEER_TIMEOUT_MS="5000")  # read once, at the library's first launch
        sys.path.insert(0, ROOT)
        import torch.distributed as dist

        import paddlepaddle_b200 as paddle
        from paddlepaddle_b200 import distributed as pdist

        pdist.init_parallel_env()
        peer = pdist.PeerGroup(slot_bytes=8 << 20)
        g = torch.Generator(device="cuda")
        g.manual_seed(100 + rank)
        # scalars: full reductions of a shard, every op, int64 exact / fp32 exact for max-min, 1e-6 for sum
        xi = torch.randint(-2**40, 2**40, (1 << 22,), device="cuda", dtype=torch.int64, generator=g)
        xf = torch.rand((1 << 22,), device="cuda", generator=g) * 2 - 1
        for it in range(7):  # odd count: both slots, several epochs
            for fn_p, fn_n in ((pdist.global_sum, pdist.global_sum), (pdist.global_max, pdist.global_max),
                               (pdist.global_min, pdist.global_min)):
                a, b = fn_p(xi, peer=peer), fn_n(xi)
                assert a.dtype == b.dtype and bool(a == b), (fn_p.__name__, a.item(), b.item())
                a, b = fn_p(xf, peer=peer), fn_n(xf)
                assert bool(a == b) if fn_p is not pdist.global_sum else abs(a.item() - b.item()) <= 1e-6 * (1 << 22), fn_p.__name__
            xi = paddle.add(xi, torch.ones_like(xi))  # new values every round: a stale slot would be caught
        # vectors: row-sharded sum(axis=0), fp32 and bf16 (float32 partials cross the link, rounded once after the
        # cross-rank sum), ragged widths; <= 64 KB of partials takes the one-shot kernel, more the two-shot
        # (reduce-scatter + all-gather) kernel
        for cols, dt in ((4096, torch.float32), (1031, torch.float32), (100003, torch.bfloat16), (8192, torch.int64),
                         (8192, torch.bfloat16), ((1 << 20) + 8, torch.float32), ((1 << 19) + 3, torch.float16),
                         (1 << 18, torch.int64)):
            rows = 512 if cols <= 200000 else 16
            if dt == torch.int64:
                x = torch.randint(-2**30, 2**30, (rows, cols), device="cuda", dtype=dt, generator=g)
            else:
                x = torch.randn((rows, cols), device="cuda", generator=g).to(dt)
            wire = torch.float32 if dt in (torch.bfloat16, torch.float16) else dt
            for it in range(3):
                a = pdist.global_sum_axis0(x, peer=peer)
                b = pdist.global_sum_axis0(x)  # NCCL on the same float32 partials, rounded once
                part = paddle.sum(x, axis=0, dtype=str(wire).split(".")[-1])
                parts = [torch.empty_like(part) for _ in range(world)]
                dist.all_gather(parts, part)
                want = parts[0].clone()
                for p_ in parts[1:]:
                    want = want + p_
                assert a.dtype == dt and b.dtype == dt
                assert torch.equal(a, want.to(dt)), (cols, dt)  # same rank order, same accumulation type: identical
                if world == 2:
                    assert torch.equal(b, a), (cols, dt, "NCCL")  # two addends: the order cannot matter
        # an unaligned `out` (a view one element into a buffer) takes the scalar path instead of faulting
        buf = torch.empty(4096 + 1, device="cuda")
        part = paddle.sum(xf.view(1024, 4096), axis=0)
        peer.slot(torch.float32, 4096).copy_(part)
        got = peer.all_reduce(4096, torch.float32, "sum", out=buf[1:])
        parts = [torch.empty_like(part) for _ in range(world)]
        dist.all_gather(parts, part)
        assert torch.equal(got, sum(parts[1:], parts[0]))
        assert peer.error() == 0
        # timing: one scalar exchange, peer kernel vs NCCL
        part = paddle.sum(xf).reshape(1)

        def timed(fn, n=200):
            for _ in range(20):
                fn()
            torch.cuda.synchronize()
            dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(n):
                fn()
            e.record()
            torch.cuda.synchronize()
            return s.elapsed_time(e) / n * 1e3

        t_nccl = timed(lambda: dist.all_reduce(part))
        t_peer = timed(lambda: peer.all_reduce(1, torch.float32, "sum"))
        assert peer.error() == 0
        # bandwidth: a 64 MiB float32 vector through the two-shot kernel vs ncclAllReduce (bytes of the vector / time)
        big_n = 16 << 20
        peer_big = pdist.PeerGroup(slot_bytes=big_n * 4)
        vec = torch.randn(big_n, device="cuda", generator=g)
        outb = torch.empty_like(vec)

        def peer_big_step():
            peer_big.slot(torch.float32, big_n).copy_(vec)
            peer_big.all_reduce(big_n, torch.float32, "sum", out=outb)

        peer_big_step()
        ref_big = vec.clone()
        dist.all_reduce(ref_big)
        assert torch.equal(outb, ref_big) if world == 2 else torch.allclose(outb, ref_big, rtol=1e-5, atol=1e-5)
        scratch = vec.clone()
        t_nccl_big = timed(lambda: dist.all_reduce(scratch), 20)
        t_copy = timed(lambda: peer_big.slot(torch.float32, big_n).copy_(vec), 20)  # advance no epoch: same slot
        t_peer_big = timed(peer_big_step, 20) - t_copy
        peer_big.close()
        # a peer that never arrives: rank 1 skips one exchange; rank 0's kernel gives up after PDB200_PEER_TIMEOUT_MS,
        # fills its output with NaN and the NEXT call on the group raises
        lonely = pdist.PeerGroup(slot_bytes=1 << 16)
        lonely.slot(torch.float32, 8).fill_(1.0)
        if rank == 0:
            res = lonely.all_reduce(8, torch.float32, "sum")
            torch.cuda.synchronize()
            assert bool(torch.isnan(res).all()), "a timed-out exchange must poison its output"
            assert lonely.error() == 1
            try:
                lonely.all_reduce(8, torch.float32, "sum")
                raise AssertionError("the call after a timeout must raise")
            except RuntimeError as exc:
                assert "timed out" in str(exc)
        lonely.close()
        peer.close()
        dist.destroy_process_group()
        q.put((rank, "ok", t_nccl, t_peer, t_nccl_big, t_peer_big))
    except Exception as exc:  # noqa: BLE001
        import traceback
        q.put((rank, "fail: " + "".join(traceback.format_exception(exc))[-1500:], 0, 0, 0, 0))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_exchange_matches_nccl():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29611, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = []
    try:
        for _ in procs:
            results.append(q.get(timeout=400))
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    for rank, status, *_ in results:
        assert status == "ok", f"rank {rank}: {status}"
    print("scalar all-reduce, us per call (rank, NCCL, peer kernel):", [(r, round(a, 1), round(b, 1)) for r, _, a, b, _, _ in results])
    print("64 MiB fp32 all-reduce, us per call and GB/s of vector bytes (rank, NCCL, peer two-shot):",
          [(r, round(c, 1), round(67.1 / c * 1e3, 1), round(d, 1), round(67.1 / d * 1e3, 1)) for r, _, _, _, c, d in results])
