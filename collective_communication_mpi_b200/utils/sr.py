"""Stochastic rounding fp32 -> bf16, bit-for-bit the arithmetic of csrc/sr.cuh, in plain torch.

Used as the oracle of the GPU tests (the kernels' updates must match it exactly) and as the CPU
statement of the optimizer step when the layers run without a GPU. See sr.cuh for why the
optimizer rounds stochastically (an SGD step on bf16 weights is usually far below one ulp).
"""
from __future__ import annotations

import torch

_M = 0xFFFFFFFF


def _hash32(x: torch.Tensor) -> torch.Tensor:
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _M
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _M
    return x ^ (x >> 16)


def bits(pair_idx: torch.Tensor, step: int, seed: int) -> torch.Tensor:
    """32 random bits (as int64 in [0, 2^32)) for each element PAIR index of ``step``."""
    p = pair_idx.to(torch.int64)
    inner = _hash32(((p >> 32) + ((step * 0x9E3779B9) & _M) + (seed & _M)) & _M)
    return _hash32((p & _M) ^ inner)


def round_bf16(x: torch.Tensor, elem_idx: torch.Tensor, step: int, seed: int) -> torch.Tensor:
    """``x`` (fp32) rounded stochastically to bf16; ``elem_idx``: each element's index in the flat
    numbering the kernel uses (element 2k and 2k+1 share draw k: low / high 16 bits)."""
    idx = elem_idx.to(torch.int64)
    r = bits(idx >> 1, step, seed)
    r16 = torch.where((idx & 1) == 0, r & 0xFFFF, (r >> 16) & 0xFFFF)
    u = x.contiguous().view(torch.int32).to(torch.int64) & _M
    finite = (u & 0x7F800000) != 0x7F800000
    u = torch.where(finite, u + r16, u)
    hi = ((u >> 16) & 0xFFFF) << 16
    hi = torch.where(hi >= 0x80000000, hi - (1 << 32), hi).to(torch.int32)
    return hi.view(torch.float32).to(torch.bfloat16)


def sgd_update(w: torch.Tensor, g: torch.Tensor, lr: float, base_index: int, step: int, seed: int) -> torch.Tensor:
    """``SR(w - lr * g)`` for a flat (or flattened) run of weights starting at ``base_index``."""
    n = w.numel()
    idx = torch.arange(base_index, base_index + n, device=w.device, dtype=torch.int64).view(w.shape)
    lr32 = torch.tensor(lr, dtype=torch.float32, device=w.device)
    return round_bf16(w.float() - lr32 * g.float(), idx, step, seed)
