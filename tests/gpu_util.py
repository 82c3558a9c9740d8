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


def half_bits_to_f32(bits, kind):
    bits = np.asarray(bits, np.uint16)
    if kind == "bf16":
        return (bits.astype(np.uint32) << 16).view(np.float32)
    return bits.view(np.float16).astype(np.float32)


def assert_half_close(got_bits, want_bits, max_ulp=1, what="", kind="bf16", atol=0.0):
    """<= max_ulp units in the last place of the 16-bit output type.  `atol` (in value units) covers
    outputs that are small differences of O(1) terms (e.g. layer_norm's y*scale + bias near zero):
    there the fp32 result itself only agrees with the CPU kernel to ~1e-5 of the terms' scale, which
    is many ulps of a tiny number."""
    got_bits, want_bits = np.asarray(got_bits).reshape(-1), np.asarray(want_bits).reshape(-1)
    d = ulp_diff_16(got_bits, want_bits)
    bad = d > max_ulp
    if atol > 0 and bad.any():
        diff = np.abs(half_bits_to_f32(got_bits, kind) - half_bits_to_f32(want_bits, kind))
        bad &= ~(diff <= atol)
    assert not bad.any(), f"{what}: {bad.sum()} elements beyond {max_ulp} ulp (max ulp distance {d[bad].max()})"
