"""Host-buffer front door: the same ops for callers whose tensors live in (pinned) host memory.

This is the end-to-end path bench.py reports as `e2e`: host -> H2D -> kernel (C-ABI) -> D2H -> host.
The output rows are independent, so the work is cut into row slabs that flow through two CUDA
streams; the H2D copy of slab i+1, the kernel of slab i and the D2H copy of slab i-1 overlap
(PCIe is full duplex, the copy engines are separate from the SMs).  The arithmetic is the same
libpdb200.so kernel as the device path; nothing is computed on the host.
"""
import torch

from . import ops

import os

_SLAB_BYTES = int(os.environ.get("PDB200_HOST_SLAB_MB", "64")) << 20
_NSTREAMS = int(os.environ.get("PDB200_HOST_STREAMS", "2"))
_state = {}


def _streams(device):
    st = _state.get(device)
    if st is None:
        st = {"streams": [torch.cuda.Stream(device=device) for _ in range(_NSTREAMS)], "bufs": {}}
        _state[device] = st
    return st


def _staging(st, slot, tag, shape, dtype, device):
    key = (slot, tag, dtype)
    n = 1
    for d in shape:
        n *= d
    buf = st["bufs"].get(key)
    if buf is None or buf.numel() < n:
        buf = torch.empty(n, dtype=dtype, device=device)
        st["bufs"][key] = buf
    return buf[:n].view(shape)


def _binary(fn_name, x, y, out=None, device=None):
    if x.is_cuda or y.is_cuda:
        raise ValueError("host.* ops take host tensors; use paddlepaddle_b200.add/multiply for device tensors")
    if x.dtype != y.dtype:
        raise TypeError("(InvalidType) the dtypes of X and Y must match")
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    oshape = torch.broadcast_shapes(x.shape, y.shape)
    if out is None:
        out = torch.empty(oshape, dtype=x.dtype, pin_memory=True)
    elif tuple(out.shape) != tuple(oshape):
        raise ValueError("(InvalidArgument) out has the wrong shape")
    # slab along the leading output dim when one operand spans it and the other either spans it too
    # or does not have it at all (the row-vector / same-shape cases); otherwise one shot.
    big, small, swap = (x, y, False) if x.shape == oshape else (y, x, True)
    lead_ok = (len(oshape) >= 1 and tuple(big.shape) == tuple(oshape) and oshape[0] > 1 and
               (tuple(small.shape) == tuple(oshape) or small.dim() < len(oshape) or small.shape[0] == 1))
    st = _streams(device)
    cur = torch.cuda.current_stream(device)
    if not lead_ok:
        dx, dy = x.to(device, non_blocking=True), y.to(device, non_blocking=True)
        out.copy_(ops._elementwise(fn_name, dx, dy), non_blocking=True)
        cur.synchronize()
        return out
    row_bytes = big[0].numel() * big.element_size()
    rows_per = max(1, min(oshape[0], _SLAB_BYTES // max(1, row_bytes)))
    small_full = tuple(small.shape) == tuple(oshape)
    d_small = None if small_full else small.to(device, non_blocking=True)
    ready = torch.cuda.Event()
    ready.record(cur)
    for k, r0 in enumerate(range(0, oshape[0], rows_per)):
        r1 = min(oshape[0], r0 + rows_per)
        slot = k % _NSTREAMS
        s = st["streams"][slot]
        with torch.cuda.stream(s):
            if k < _NSTREAMS:
                s.wait_event(ready)
            shp = (r1 - r0,) + tuple(oshape[1:])
            d_big = _staging(st, slot, "big", shp, x.dtype, device)
            d_out = _staging(st, slot, "out", shp, x.dtype, device)
            d_big.copy_(big[r0:r1], non_blocking=True)
            if small_full:
                d_sm = _staging(st, slot, "small", shp, x.dtype, device)
                d_sm.copy_(small[r0:r1], non_blocking=True)
            else:
                d_sm = d_small
            a, b = (d_sm, d_big) if swap else (d_big, d_sm)
            ops._elementwise(fn_name, a, b, out=d_out)  # launches on stream s (torch's current stream)
            out[r0:r1].copy_(d_out, non_blocking=True)
    for s in st["streams"]:
        s.synchronize()
    return out


def add(x, y, out=None, device=None):
    """paddle.add for host tensors (pinned memory recommended); returns a host tensor."""
    return _binary("pdb200_add", x, y, out, device)


def multiply(x, y, out=None, device=None):
    return _binary("pdb200_multiply", x, y, out, device)
