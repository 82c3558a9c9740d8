"""paddle.distributed subset used by the hot path: one process per GPU, NCCL over NVLink.

Mirrors python/paddle/distributed/parallel.py:978 (init_parallel_env),
communication/all_reduce.py + stream/all_reduce.py:39-55 (all_reduce -> ProcessGroupNCCL::AllReduce,
paddle/fluid/distributed/collective/process_group_nccl.cc:267-297 -> ncclAllReduce,
paddle/phi/core/distributed/nccl_comm_context.cc:199-205).  torch.distributed is the plumbing.

Only two things on the hot path ever cross GPUs:
  * nothing, for row-independent ops: `shard_rows` gives each rank its contiguous block of rows;
  * one scalar (or one (value, index) pair) per rank for full-tensor reductions: `global_*` reduce
    the local shard with the CUDA kernels and finish with a single AllReduce on the native dtype.
"""
import os

import torch
import torch.distributed as dist

from . import ops


class ReduceOp:
    SUM = dist.ReduceOp.SUM
    MAX = dist.ReduceOp.MAX
    MIN = dist.ReduceOp.MIN
    PROD = dist.ReduceOp.PRODUCT


def init_parallel_env(backend=None):
    """One process per GPU; rendezvous from MASTER_ADDR/MASTER_PORT/RANK/WORLD_SIZE (torchrun)."""
    if dist.is_initialized():
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group(backend)


def get_rank():
    return dist.get_rank() if dist.is_initialized() else 0


def get_world_size():
    return dist.get_world_size() if dist.is_initialized() else 1


def all_reduce(tensor, op=ReduceOp.SUM, group=None, sync_op=True):
    """In-place all-reduce (paddle.distributed.all_reduce).  On the calc stream, like
    use_calc_stream=True (process_group_nccl.cc:930)."""
    if get_world_size() == 1:
        return None
    work = dist.all_reduce(tensor, op=op, group=group, async_op=not sync_op)
    return work


def shard_rows(n_rows, rank=None, world=None):
    """Contiguous block [begin, end) of rows owned by `rank`; the first n_rows % world ranks get one extra."""
    rank = get_rank() if rank is None else rank
    world = get_world_size() if world is None else world
    base, extra = divmod(int(n_rows), world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def combine_argmax(local_val, local_idx, shard_begin):
    """(value, global index) pairs -> global argmax with the lowest index among ties:
    AllReduce(MAX) on the value, then AllReduce(MIN) on the index masked to ranks holding that value."""
    gval = local_val.clone()
    all_reduce(gval, ReduceOp.MAX)
    big = torch.iinfo(torch.int64).max
    cand = torch.where(local_val == gval, local_idx.to(torch.int64) + int(shard_begin),
                       torch.full_like(local_idx, big, dtype=torch.int64))
    all_reduce(cand, ReduceOp.MIN)
    return gval, cand


def global_sum(x_shard, dtype=None):
    """sum over a tensor whose flat elements are sharded across ranks."""
    part = ops.sum(x_shard, dtype=dtype).reshape(1)
    all_reduce(part, ReduceOp.SUM)
    return part.reshape(())


def global_max(x_shard):
    part = ops.max(x_shard).reshape(1)
    all_reduce(part, ReduceOp.MAX)
    return part.reshape(())


def global_min(x_shard):
    part = ops.min(x_shard).reshape(1)
    all_reduce(part, ReduceOp.MIN)
    return part.reshape(())


def global_prod(x_shard):
    part = ops.prod(x_shard).reshape(1)
    all_reduce(part, ReduceOp.PROD)
    return part.reshape(())


def global_mean(x_shard, total_numel):
    """mean = global sum / total element count (sum then scale)."""
    s = global_sum(x_shard)
    return s / total_numel if s.is_floating_point() else torch.div(s, total_numel, rounding_mode="trunc")


def global_argmax(x_shard, shard_begin):
    """argmax over the flattened global tensor from per-rank shards."""
    idx = ops.argmax(x_shard).reshape(1)
    val = x_shard.reshape(-1)[idx]
    _, gidx = combine_argmax(val, idx, shard_begin)
    return gidx.reshape(())
