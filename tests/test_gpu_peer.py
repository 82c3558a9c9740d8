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
                          LOCAL_RANK=str(rank))
        sys.path.insert(0, ROOT)
        import torch.distributed as dist

        import paddlepaddle_b200 as paddle
        from paddlepaddle_b200 import distributed as pdist

        pdist.init_parallel_env()
        peer = pdist.PeerGroup(slot_bytes=1 << 20)
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
        # vectors: row-sharded sum(axis=0), fp32 and bf16 (accumulated in float, rounded once), ragged width
        for cols, dt in ((4096, torch.float32), (1031, torch.float32), (100003, torch.bfloat16), (8192, torch.int64)):
            if dt == torch.int64:
                x = torch.randint(-2**30, 2**30, (512, cols), device="cuda", dtype=dt, generator=g)
            else:
                x = torch.randn((512, cols), device="cuda", generator=g).to(dt)
            for it in range(3):
                a = pdist.global_sum_axis0(x, peer=peer)
                part = paddle.sum(x, axis=0)
                parts = [torch.empty_like(part) for _ in range(world)]
                dist.all_gather(parts, part)
                want = parts[0].float() if dt != torch.int64 else parts[0].clone()
                for p_ in parts[1:]:
                    want = want + (p_.float() if dt != torch.int64 else p_)
                if dt == torch.int64:
                    assert torch.equal(a, want)
                else:
                    assert torch.equal(a, want.to(dt)), (cols, dt)  # same rank order, same float accumulation: identical
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
        peer.close()
        dist.destroy_process_group()
        q.put((rank, "ok", t_nccl, t_peer))
    except Exception as exc:  # noqa: BLE001
        import traceback
        q.put((rank, "fail: " + "".join(traceback.format_exception(exc))[-1500:], 0, 0))


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
            results.append(q.get(timeout=240))
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    for rank, status, t_nccl, t_peer in results:
        assert status == "ok", f"rank {rank}: {status}"
    print("scalar all-reduce, us per call (rank, NCCL, peer kernel):", [(r, round(a, 1), round(b, 1)) for r, _, a, b in results])
