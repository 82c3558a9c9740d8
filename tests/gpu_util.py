"""Helpers for the GPU parity tests: numpy <-> device tensors, ulp distances."""
import numpy as np
import torch

import oracle_lib as O


def to_dev(a, half=None):
    """numpy -> cuda tensor.  half='bf16'|'fp16' reinterprets a uint16 bit-pattern array."""
    a = np.ascontiguousarray(a)
    if half == "bf16":
        return torch.from_numpy(a.view(np.int16).copy()).cuda().view(torch.bfloat16)
    if half == "fp16":
        return torch.from_numpy(a.view(np.int16).copy()).cuda().view(torch.float16)
    return torch.from_numpy(a.copy()).cuda()


def to_np(t):
    """cuda tensor -> numpy; bf16/fp16 come back as uint16 bit patterns."""
    t = t.detach()
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.view(torch.int16).cpu().numpy().view(np.uint16)
    return t.cpu().numpy()


def ulp_diff_16(a_bits, b_bits):
    """Distance in units of last place between two arrays of 16-bit float bit patterns."""
    def key(u):
        u = u.astype(np.int32)
        return np.where(u & 0x8000, 0x8000 - u, u)  # monotone integer key, -0 == +0
    return np.abs(key(a_bits) - key(b_bits))


def assert_half_close(got_bits, want_bits, max_ulp=1, what=""):
    d = ulp_diff_16(got_bits, want_bits)
    assert d.max(initial=0) <= max_ulp, f"{what}: max ulp distance {d.max()} (> {max_ulp}), mismatches {(d > max_ulp).sum()}"
