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
import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib as L
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


class PeerGroup:
    """All-reduce of small results over NVLink peer memory, without NCCL (csrc/peer.cu, include/pdb200.h).

    Every rank owns a window of device memory (flags + two slots) that all other ranks of the node map through CUDA
    IPC.  `slot(dtype, shape)` hands out the current slot as a tensor — pass it as the output of the local reduction —
    and `all_reduce(n, op)` runs the one exchange kernel: publish this rank's epoch to all peers, wait for theirs,
    combine all slots in rank order.  No host synchronisation, no staging copy; results are bit-identical on all ranks.
    Set-up (the constructor) is collective: it exchanges the 64-byte IPC handles with one all_gather."""

    _OPS = {"sum": 0, "max": 1, "min": 2, "prod": 3}

    def __init__(self, slot_bytes=1 << 20):
        if not dist.is_initialized():
            raise RuntimeError("PeerGroup needs an initialised process group (init_parallel_env)")
        self.lib = L.load()
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.slot_bytes = int(slot_bytes)
        self.epoch = 0
        self._window = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        L.check(self.lib.pdb200_peer_window_create(self.slot_bytes, ctypes.byref(self._window), handle), self.lib)
        mine = torch.tensor(list(handle), dtype=torch.uint8, device="cuda")
        handles = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(handles, mine)
        self._mapped = (ctypes.c_void_p * self.world)()
        for r, h in enumerate(handles):
            if r == self.rank:
                self._mapped[r] = self._window.value
                continue
            raw = (ctypes.c_ubyte * 64)(*h.cpu().tolist())
            ptr = ctypes.c_void_p()
            L.check(self.lib.pdb200_peer_window_open(raw, ctypes.byref(ptr)), self.lib)
            self._mapped[r] = ptr.value
        dist.barrier()  # every rank has mapped every window before the first exchange

    def slot(self, dtype, numel):
        """The slot of the NEXT all_reduce as a flat tensor of `numel` elements (write the local partial result here)."""
        esize = torch.empty((), dtype=dtype).element_size()
        if numel * esize > self.slot_bytes:
            raise ValueError(f"(InvalidArgument) {numel} x {dtype} does not fit the {self.slot_bytes}-byte slot")
        ptr = self.lib.pdb200_peer_slot(self._window, self.slot_bytes, self.epoch + 1)
        return _tensor_from_ptr(ptr, numel, dtype)

    def all_reduce(self, numel, dtype, op="sum", out=None, slot_dtype=None):
        """Combine the `numel` elements every rank wrote into its slot; returns a new tensor (or fills `out`).
        `slot_dtype`: what the slots hold when it is not `dtype` (float32 partials of a bfloat16 / float16 sum — the
        cross-rank sum is then rounded once).  Raises if an EARLIER exchange of this group timed out (checked without
        synchronising: the kernel raises a host-visible word); a timed-out exchange leaves NaN / all-ones in its output."""
        self.epoch += 1
        if out is None:
            out = torch.empty(numel, dtype=dtype, device="cuda")
        L.check(self.lib.pdb200_peer_all_reduce(self._OPS[op], self._mapped, self.world, self.rank, self.slot_bytes, self.epoch,
                                                numel, ops._dtype_code(slot_dtype or dtype), ops._dtype_code(dtype),
                                                ctypes.c_void_p(out.data_ptr()),
                                                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), self.lib)
        return out

    def error(self):
        """Non-zero if an exchange timed out waiting for a peer (synchronises the device)."""
        return self.lib.pdb200_peer_error(self._window)

    def close(self):
        torch.cuda.synchronize()
        dist.barrier()
        for r in range(self.world):
            if r != self.rank and self._mapped[r]:
                self.lib.pdb200_peer_window_close(ctypes.c_void_p(self._mapped[r]))
        dist.barrier()
        self.lib.pdb200_peer_window_destroy(self._window)
        self._window = ctypes.c_void_p()


class _ForeignMemory:
    """__cuda_array_interface__ view of device memory owned by libpdb200.so (a peer window slot)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def _tensor_from_ptr(ptr, numel, dtype):
    esize = torch.empty((), dtype=dtype).element_size()
    return torch.as_tensor(_ForeignMemory(ptr, numel * esize), device="cuda").view(dtype)


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


_LOCAL = {"sum": (L.REDUCE_SUM, ReduceOp.SUM), "max": (L.REDUCE_MAX, ReduceOp.MAX), "min": (L.REDUCE_MIN, ReduceOp.MIN),
          "prod": (L.REDUCE_PROD, ReduceOp.PROD)}


def _global(op_name, x_shard, peer, axis=None, out_dtype=None):
    """Local reduction, then the exchange: through `peer` (a PeerGroup — the reduction kernel writes its result straight
    into this rank's peer-visible slot and ONE kernel combines all ranks' slots over NVLink) or, without one, an NCCL
    AllReduce.  Sums and products of bfloat16 / float16 travel as the float32 partials the kernels accumulate in and
    are rounded ONCE, after the cross-rank combine ("1 ulp of the output type, accumulated in fp32" holds for the
    global result, not per rank)."""
    code, nccl_op = _LOCAL[op_name]
    dt = out_dtype or x_shard.dtype
    if op_name == "sum" and out_dtype is None and x_shard.dtype in (torch.bool, torch.int32):
        dt = torch.int64  # SumRawInferMeta
    # the dtype that crosses GPUs: float32 for 16-bit float sums (sum_raw has the bf16/fp16 -> fp32 kernel)
    wire = torch.float32 if (op_name == "sum" and dt in (torch.bfloat16, torch.float16)) else dt
    if get_world_size() == 1:
        return ops._reduce(code, x_shard, axis, False, dt).reshape(-1)
    if peer is not None:
        n = 1 if axis is None else x_shard.shape[-1]
        ops._reduce(code, x_shard, axis, False, wire, out=peer.slot(wire, n))
        return peer.all_reduce(n, dt, op_name, slot_dtype=wire)
    part = ops._reduce(code, x_shard, axis, False, wire).reshape(-1)
    all_reduce(part, nccl_op)
    return part if wire == dt else ops.cast(part, dt)


def global_sum(x_shard, dtype=None, peer=None):
    """sum over a tensor whose flat elements are sharded across ranks."""
    return _global("sum", x_shard, peer, out_dtype=ops._as_torch_dtype(dtype)).reshape(())


def global_max(x_shard, peer=None):
    return _global("max", x_shard, peer).reshape(())


def global_min(x_shard, peer=None):
    return _global("min", x_shard, peer).reshape(())


def global_prod(x_shard, peer=None):
    return _global("prod", x_shard, peer).reshape(())


def global_sum_axis0(x_shard, peer=None):
    """sum(axis=0) of a matrix whose ROWS are sharded across ranks: local column sums, then the exchange of the
    [cols] vector (SURVEY.md §8e last row / §8f rank 4)."""
    return _global("sum", x_shard.reshape(-1, x_shard.shape[-1]), peer, axis=0)


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
