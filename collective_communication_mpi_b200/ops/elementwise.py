"""Fused elementwise ops of the transformer-FC block (csrc/elementwise.cu): the gate that stands in
for attention (``v * sigmoid(q * k)``), tanh-GELU, and the MSE loss with its gradient -- each one
pass over HBM, with hand-written backward kernels instead of the eager op chains autograd would
record. The reference has no compute (func_impl.py stops at the collectives), so there is nothing
to mirror; the ``*_reference`` functions below state the math in plain PyTorch and are the oracle
of the tests.

CUDA tensors always go to the kernels (a missing library is an error, not a fallback). CPU tensors
-- the CPU test-suite, documentation examples -- run the reference formulas.

The models use these ops by default (``CCMPI_FUSED_ELEMENTWISE=0`` switches back to eager PyTorch op
chains, which is what the NCCL + cuBLAS baseline arm runs). Measured on a B200
(profiles/r2_extras_bench_1gpu.jsonl): gate fwd+bwd 4.5x, MSE loss+grad 7.7x the eager chains.
"""
from __future__ import annotations

import math
import os
from typing import Tuple

import torch

from .. import _native as N

_C = math.sqrt(2.0 / math.pi)
_A = 0.044715


def enabled() -> bool:
    return os.environ.get("CCMPI_FUSED_ELEMENTWISE", "1") != "0"


def _stream(stream=None) -> int:
    return (stream if stream is not None else torch.cuda.current_stream()).cuda_stream


# -------------------------------------------------------------------------------------------------
# the math, in fp32 PyTorch (oracle; also the CPU path)
# -------------------------------------------------------------------------------------------------
def gate_reference(qkv: torch.Tensor) -> torch.Tensor:
    q, k, v = qkv.float().chunk(3, dim=-1)
    return (v * torch.sigmoid(q * k)).to(qkv.dtype)


def gate_backward_reference(da: torch.Tensor, qkv: torch.Tensor) -> torch.Tensor:
    q, k, v = qkv.float().chunk(3, dim=-1)
    g = da.float()
    s = torch.sigmoid(q * k)
    dz = g * v * s * (1.0 - s)
    return torch.cat([dz * k, dz * q, g * s], dim=-1).to(qkv.dtype)


def gelu_reference(u: torch.Tensor) -> torch.Tensor:
    x = u.float()
    return (0.5 * x * (1.0 + torch.tanh(_C * (x + _A * x ** 3)))).to(u.dtype)


def gelu_backward_reference(dy: torch.Tensor, u: torch.Tensor) -> torch.Tensor:
    x = u.float()
    th = torch.tanh(_C * (x + _A * x ** 3))
    dth = (1.0 - th * th) * _C * (1.0 + 3.0 * _A * x * x)
    return (dy.float() * (0.5 * (1.0 + th) + 0.5 * x * dth)).to(u.dtype)


def mse_reference(y: torch.Tensor, t: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    d = y.float() - t.float()
    return (d * d).mean(), (d * (2.0 / d.numel())).to(y.dtype)


# -------------------------------------------------------------------------------------------------
# kernels
# -------------------------------------------------------------------------------------------------
def _rows(x: torch.Tensor) -> torch.Tensor:
    if x.dtype != torch.bfloat16:
        raise TypeError(f"fused elementwise kernels take bfloat16, got {x.dtype}")
    x2 = x.reshape(-1, x.shape[-1])
    if x2.stride(1) != 1:
        x2 = x2.contiguous()
    return x2


def gate_forward(qkv: torch.Tensor, stream=None) -> torch.Tensor:
    if not qkv.is_cuda:
        return gate_reference(qkv)
    x = _rows(qkv)
    T, H3 = x.shape
    if H3 % 3:
        raise ValueError("last dimension must hold q | k | v side by side")
    H = H3 // 3
    out = torch.empty((T, H), dtype=qkv.dtype, device=qkv.device)
    N.check(N.device_lib().ccmpi_gate_fwd(x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), T, H, _stream(stream)),
            "ccmpi_gate_fwd")
    N.count_launch()
    return out.view(*qkv.shape[:-1], H)


def gate_backward(da: torch.Tensor, qkv: torch.Tensor, stream=None) -> torch.Tensor:
    if not qkv.is_cuda:
        return gate_backward_reference(da, qkv)
    x, g = _rows(qkv), _rows(da)
    T, H3 = x.shape
    H = H3 // 3
    out = torch.empty((T, H3), dtype=qkv.dtype, device=qkv.device)
    N.check(N.device_lib().ccmpi_gate_bwd(g.data_ptr(), g.stride(0), x.data_ptr(), x.stride(0), out.data_ptr(),
                                          out.stride(0), T, H, _stream(stream)), "ccmpi_gate_bwd")
    N.count_launch()
    return out.view(qkv.shape)


def gelu_forward(u: torch.Tensor, stream=None) -> torch.Tensor:
    if not u.is_cuda:
        return gelu_reference(u)
    x = u if u.is_contiguous() else u.contiguous()
    if x.dtype != torch.bfloat16:
        raise TypeError(f"fused elementwise kernels take bfloat16, got {x.dtype}")
    out = torch.empty_like(x)
    N.check(N.device_lib().ccmpi_gelu_fwd(x.data_ptr(), out.data_ptr(), x.numel(), _stream(stream)), "ccmpi_gelu_fwd")
    N.count_launch()
    return out


def gelu_backward(dy: torch.Tensor, u: torch.Tensor, stream=None) -> torch.Tensor:
    if not u.is_cuda:
        return gelu_backward_reference(dy, u)
    x = u if u.is_contiguous() else u.contiguous()
    g = dy if dy.is_contiguous() else dy.contiguous()
    out = torch.empty_like(x)
    N.check(N.device_lib().ccmpi_gelu_bwd(g.data_ptr(), x.data_ptr(), out.data_ptr(), x.numel(), _stream(stream)),
            "ccmpi_gelu_bwd")
    N.count_launch()
    return out


_mse_ws = {}


def mse_loss_and_grad(y: torch.Tensor, t: torch.Tensor, stream=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(mean((y - t)^2), d loss / d y)`` in one pass; the loss is a 0-d fp32 tensor on ``y``'s device."""
    if not y.is_cuda:
        return mse_reference(y, t)
    if y.dtype != torch.bfloat16 or t.dtype != torch.bfloat16 or y.shape != t.shape:
        raise TypeError("mse_loss_and_grad takes two bfloat16 tensors of the same shape")
    yc = y if y.is_contiguous() else y.contiguous()
    tc = t if t.is_contiguous() else t.contiguous()
    key = (y.device.index, _stream(stream))
    ws = _mse_ws.get(key)
    if ws is None:      # zeroed once; the kernel leaves it ready for the next call on the same stream
        ws = torch.zeros(N.device_lib().ccmpi_mse_workspace_bytes(), dtype=torch.uint8, device=y.device)
        _mse_ws[key] = ws
    dy = torch.empty_like(yc)
    loss = torch.empty((), dtype=torch.float32, device=y.device)
    N.check(N.device_lib().ccmpi_mse_fwd_bwd(yc.data_ptr(), tc.data_ptr(), dy.data_ptr(), yc.numel(), ws.data_ptr(),
                                             loss.data_ptr(), _stream(stream)), "ccmpi_mse_fwd_bwd")
    N.count_launch()
    return loss, dy


def sgd_sr_(w: torch.Tensor, g: torch.Tensor, lr: float, *, base_index: int = 0, step: int = 0, seed: int = 0, stream=None):
    """``w = SR(w - lr * g)`` in place on contiguous bf16 tensors: the optimizer pass for gradients that came
    out of a collective, with stochastic rounding (csrc/sr.cuh) so that steps below one bf16 ulp are not
    lost. ``base_index``: index of ``w``'s first element in the model's flat parameter numbering."""
    if not w.is_cuda:
        from ..utils import sr
        w.copy_(sr.sgd_update(w.reshape(-1), g.reshape(-1), lr, base_index, step & 0xFFFFFFFF, seed).view(w.shape))
        return w
    if w.dtype != torch.bfloat16 or g.dtype != torch.bfloat16 or not w.is_contiguous() or not g.is_contiguous() or w.shape != g.shape:
        raise TypeError("sgd_sr_ takes two contiguous bfloat16 tensors of the same shape")
    N.check(N.device_lib().ccmpi_sgd_sr_bf16(w.data_ptr(), g.data_ptr(), w.numel(), float(lr), int(base_index),
                                             int(step) & 0xFFFFFFFF, int(seed) & 0xFFFFFFFF, _stream(stream)), "ccmpi_sgd_sr_bf16")
    N.count_launch()
    return w


# -------------------------------------------------------------------------------------------------
# autograd
# -------------------------------------------------------------------------------------------------
class _Gate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv):
        ctx.save_for_backward(qkv)
        return gate_forward(qkv)

    @staticmethod
    def backward(ctx, da):
        (qkv,) = ctx.saved_tensors
        return gate_backward(da, qkv)


class _Gelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u):
        ctx.save_for_backward(u)
        return gelu_forward(u)

    @staticmethod
    def backward(ctx, dy):
        (u,) = ctx.saved_tensors
        return gelu_backward(dy, u)


def gate(qkv: torch.Tensor) -> torch.Tensor:
    """``v * sigmoid(q * k)`` for ``qkv = q | k | v`` along the last axis; differentiable."""
    return _Gate.apply(qkv)


def gelu(u: torch.Tensor) -> torch.Tensor:
    """tanh-approximated GELU; differentiable."""
    return _Gelu.apply(u)
