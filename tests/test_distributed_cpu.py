"""world_size-2 gloo run of the host-side multi-GPU logic (no GPU needed): shard partitioning and the
exchange step of full-tensor reductions.  The per-rank partials come from the CPU oracle here (the
product kernels only run on CUDA); the combine step is the product code in paddlepaddle_b200.distributed."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import oracle_lib as O
    from paddlepaddle_b200 import distributed as pd
    pd.init_parallel_env("gloo")
    port_oracle = O.Oracle("port")
    n = 100003
    rng = np.random.default_rng(123)  # every rank regenerates the same global tensor, keeps its shard
    xi = rng.integers(-2**40, 2**40, n, dtype=np.int64)
    xf = rng.integers(0, 50, n).astype(np.float32)  # ties across shards
    b, e = pd.shard_rows(n)
    res = {"shard": (b, e)}
    part = torch.from_numpy(port_oracle.reduce("sum", xi[b:e]).reshape(1).copy())
    pd.all_reduce(part, pd.ReduceOp.SUM)
    res["sum"] = int(part)
    part = torch.from_numpy(port_oracle.reduce("max", xi[b:e]).reshape(1).copy())
    pd.all_reduce(part, pd.ReduceOp.MAX)
    res["max"] = int(part)
    li = port_oracle.argmax(xf[b:e]).reshape(1)
    lv = xf[b:e][li]
    gv, gi = pd.combine_argmax(torch.from_numpy(lv.copy()), torch.from_numpy(li.copy()), b)
    res["argmax"] = int(gi)
    res["want"] = (int(port_oracle.reduce("sum", xi)), int(xi.max()), int(port_oracle.argmax(xf)))
    q.put((rank, res))
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_two_rank_gloo_exchange():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out[0]["shard"] == (0, 50002) and out[1]["shard"] == (50002, 100003)
    for r in range(world):
        assert (out[r]["sum"], out[r]["max"], out[r]["argmax"]) == out[r]["want"]


def test_shard_rows_cover_exactly():
    from paddlepaddle_b200 import distributed as pd
    for n in (65536, 7, 100003, 0):
        for world in (1, 2, 4, 8):
            spans = [pd.shard_rows(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
