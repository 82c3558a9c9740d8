"""The NVLink exchange kernels (csrc/peer.cu) on ONE GPU: every rank of a 2-, 4- or 8-rank group as one cooperative grid
(pdb200_peer_all_reduce_emulated, blockIdx.y = rank), with all windows in this GPU's memory.  The device code is the
code the multi-GPU path runs — flags, epochs, alternating slots and result regions, the one-shot read-everything form,
the two-shot reduce-scatter + all-gather form with its ticket counter and second flag — so a box with a single GPU (the
driver's GPU test tier) still exercises the protocol; tests/test_gpu_peer.py runs the real thing on two GPUs.
Expected values: the partials summed in rank order in the slot type, rounded once to the output type."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

from paddlepaddle_b200 import _lib as L  # noqa: E402
from paddlepaddle_b200 import ops  # noqa: E402
from paddlepaddle_b200.distributed import _tensor_from_ptr  # noqa: E402


class _Group:
    def __init__(self, world, slot_bytes):
        self.lib = L.load()
        self.world, self.slot_bytes, self.epoch = world, slot_bytes, 0
        self.windows = (ctypes.c_void_p * world)()
        for r in range(world):
            w = ctypes.c_void_p()
            handle = (ctypes.c_ubyte * 64)()
            L.check(self.lib.pdb200_peer_window_create(slot_bytes, ctypes.byref(w), handle), self.lib)
            self.windows[r] = w.value

    def slot(self, rank, dtype, n):
        ptr = self.lib.pdb200_peer_slot(ctypes.c_void_p(self.windows[rank]), self.slot_bytes, self.epoch + 1)
        return _tensor_from_ptr(ptr, n, dtype)

    def all_reduce(self, op, n, slot_dtype, out_dtype):
        self.epoch += 1
        outs = [torch.empty(n, dtype=out_dtype, device="cuda") for _ in range(self.world)]
        ptrs = (ctypes.c_void_p * self.world)(*[o.data_ptr() for o in outs])
        L.check(self.lib.pdb200_peer_all_reduce_emulated({"sum": 0, "max": 1}[op], self.windows, self.world, self.slot_bytes,
                                                         self.epoch, n, ops._dtype_code(slot_dtype), ops._dtype_code(out_dtype),
                                                         ptrs, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), self.lib)
        return outs

    def error(self):
        return [self.lib.pdb200_peer_error(ctypes.c_void_p(w)) for w in self.windows]

    def close(self):
        torch.cuda.synchronize()
        for w in self.windows:
            self.lib.pdb200_peer_window_destroy(ctypes.c_void_p(w))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_exchange_protocol_all_ranks_in_one_grid(world):
    g = torch.Generator(device="cuda")
    g.manual_seed(world)
    grp = _Group(world, slot_bytes=8 << 20)
    try:
        cases = [(1, torch.float32, torch.float32, "sum"), (1, torch.int64, torch.int64, "max"),
                 (4096, torch.float32, torch.float32, "sum"),            # 16 KB: one-shot
                 (1031, torch.float32, torch.bfloat16, "sum"),           # ragged, rounded once at the end
                 (8192, torch.int64, torch.int64, "sum"),                # 64 KB: still one-shot
                 (100003, torch.float32, torch.bfloat16, "sum"),         # 400 KB: two-shot, ragged tail
                 ((1 << 20) + 8, torch.float32, torch.float32, "sum"),   # 4 MiB: two-shot, several trips per thread
                 (1 << 18, torch.int64, torch.int64, "max")]
        for rep in range(3):  # three epochs per case: both slots / result regions, and a reuse
            for n, sdt, odt, op in cases:
                parts = []
                for r in range(world):
                    if sdt == torch.int64:
                        p = torch.randint(-2**40, 2**40, (n,), device="cuda", dtype=sdt, generator=g)
                    else:
                        p = torch.randn(n, device="cuda", generator=g)
                    grp.slot(r, sdt, n).copy_(p)
                    parts.append(p)
                outs = grp.all_reduce(op, n, sdt, odt)
                want = parts[0].clone()
                for p in parts[1:]:
                    want = want + p if op == "sum" else torch.maximum(want, p)
                want = want.to(odt)
                for r, o in enumerate(outs):
                    assert torch.equal(o, want), (world, n, sdt, odt, op, rep, r)
        assert grp.error() == [0] * world
    finally:
        grp.close()
