"""Helpers of the GPU parity tests: logical views of the native plan's arena (the "planes" layout of
open_universe_b200/csrc/uv_common.cuh) read / written through the C ABI (uv_arena_read / uv_arena_write)."""
import ctypes as C

import torch

from open_universe_b200 import _abi, _native


def _dtype(t):
    return torch.bfloat16 if t.dtype == _abi.DT_BF16 else torch.float32


def read_planes(plan, t, B, L):
    """Planes tensor `t` of the plan's arena -> logical [B, C, L] fp32 (CPU)."""
    es = 2 if t.dtype == _abi.DT_BF16 else 4
    n = t.split * t.nch * t.P * 8
    raw = plan.arena_read(t.off, n * es)
    v = torch.frombuffer(bytearray(raw), dtype=_dtype(t)).float().view(t.split, t.nch, t.P, 8)
    rows = (L + t.split - 1) // t.split
    out = torch.zeros(B, t.nch * 8, rows * t.split)
    for b in range(B):
        blk = v[:, :, t.G + b * t.Lh : t.G + b * t.Lh + rows, :]   # [split, nch, rows, 8]
        out[b] = blk.permute(1, 3, 2, 0).reshape(t.nch * 8, rows * t.split)
    return out[:, :, :L]


def write_planes(plan, t, x):
    """Logical [B, C, L] tensor -> planes tensor `t` (plain, split 1) of the plan's arena; padding rows stay zero."""
    assert t.split == 1
    B, Cc, L = x.shape
    assert Cc == t.nch * 8
    v = torch.zeros(t.nch, t.P, 8, dtype=_dtype(t))
    for b in range(B):
        v[:, t.G + b * t.Lh : t.G + b * t.Lh + L, :] = x[b].view(t.nch, 8, L).permute(0, 2, 1).to(v.dtype)
    raw = v.contiguous().view(torch.uint8).numpy().tobytes()
    buf = C.create_string_buffer(raw, len(raw))
    _native._check(_native.lib().uv_arena_write(plan.handle, t.off, buf, len(raw)), "uv_arena_write")


def read_f32(plan, off, n):
    return torch.frombuffer(bytearray(plan.arena_read(off, 4 * n)), dtype=torch.float32).clone()
