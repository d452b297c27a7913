"""Tensor-parallel FC layers in the classic Megatron layout: activations are REPLICATED across the
model-parallel group outside a column/row pair and feature-sharded inside it.

``MegatronColumnParallelLinear`` (weight ``[in, out/n]``)
    fwd  ``y_loc = x @ W``                      plain GEMM, no communication
    bwd  ``dx = all_reduce(dy_loc @ W^T)``      GEMM -> all-reduce (every rank needs the full dx)
         ``dW = x^T @ dy_loc``

``MegatronRowParallelLinear`` (weight ``[in/n, out]``)
    fwd  ``y = all_reduce(x_loc @ W)``          GEMM -> all-reduce
    bwd  ``dx_loc = dy @ W^T``, ``dW = x_loc^T @ dy``   plain GEMMs

Compared with the reference's layout (``tp_layers.py``: activations stay feature-sharded, one gather and
one scatter per layer, func_impl.py:76-187) this moves 2 x [M, hidden] per pair through all-reduces and
keeps the residual stream whole on every rank. The reference's ``get_info`` partitions weights exactly
this way (column-parallel fc_q/k/v, row-parallel fc_o, func_impl.py:65-70), so both layouts share shards.

The all-reduce is fused into the producing GEMM (``ops.fused.GemmAllReduce``, ArPolicy in
csrc/fused_policies.cuh); the layers are checked against a dense model on CPU ranks and on 4 / 8 B200s
(tests/test_gpu_seq.py::test_megatron_layout*).
"""
from __future__ import annotations

import torch
from torch import nn

from .tp_layers import _backend, _dw_last, _init_weight, _weight_grad, set_tokens


class _ColFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, layer):
        ctx.layer = layer
        ctx.save_for_backward(x, weight)
        return layer._gemm(x, weight, b_kn=True, num_sms=layer.num_sms)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        layer, dy = ctx.layer, dy.contiguous()
        fused_sgd = _dw_last(layer)       # dW's kernel (or the reduction it triggers) rewrites W: dx first
        dw = None if fused_sgd else _weight_grad(layer, x, dy)
        if layer.mp == 1:
            dx = layer._gemm(dy, weight, num_sms=layer.num_sms)
        else:
            # the result lives in a double-buffered symmetric buffer: hand autograd its own copy
            dx = layer.ar(dy, weight, b_kn=False, num_sms=layer.num_sms).clone()
        if fused_sgd:
            dw = _weight_grad(layer, x, dy)          # rewrites W: after dx has read it
        return dx, dw, None


class _RowFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, layer):
        ctx.layer = layer
        ctx.save_for_backward(x, weight)
        if layer.mp == 1:
            return layer._gemm(x, weight, b_kn=True, num_sms=layer.num_sms)
        return layer.ar(x, weight, b_kn=True, num_sms=layer.num_sms).clone()

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        layer, dy = ctx.layer, dy.contiguous()
        dx = layer._gemm(dy, weight, num_sms=layer.num_sms)
        dw = _weight_grad(layer, x, dy)
        return dx, dw, None


class MegatronColumnParallelLinear(nn.Module):
    """``[M, in]`` (replicated) ``-> [M, out/n]``."""

    def __init__(self, mp_comm, tokens: int, in_features: int, out_features: int, *, grad_dtype=torch.bfloat16, seed: int = 0):
        super().__init__()
        self.comm, self.mp = mp_comm, mp_comm.Get_size()
        self.M, self.K, self.N = tokens, in_features, out_features // self.mp
        self.grad_dtype = grad_dtype
        dev, ops, self._gemm, self._matmul_tn, self._gemm_sgd = _backend()
        self.sgd_lr = None
        self.weight = _init_weight(in_features, self.N, in_features, dev, seed * 1000 + mp_comm.Get_rank())
        self._ops = ops
        self._build_ops()
        self.num_sms = 0

    def _build_ops(self):
        self.ar = self._ops.GemmAllReduce(self.comm, self.M, self.K) if self.mp > 1 else None      # backward dx

    def set_tokens(self, tokens: int) -> None:
        """Rebuild the fused op for another token count (COLLECTIVE; ``tp_layers.set_tokens``)."""
        set_tokens(self, tokens)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _ColFn.apply(x.reshape(self.M, -1), self.weight, self)


class MegatronRowParallelLinear(nn.Module):
    """``[M, in/n] -> [M, out]`` (replicated): partial products are all-reduced."""

    def __init__(self, mp_comm, tokens: int, in_features: int, out_features: int, *, grad_dtype=torch.bfloat16, seed: int = 0):
        super().__init__()
        self.comm, self.mp = mp_comm, mp_comm.Get_size()
        self.M, self.K, self.N = tokens, in_features // self.mp, out_features
        self.grad_dtype = grad_dtype
        dev, ops, self._gemm, self._matmul_tn, self._gemm_sgd = _backend()
        self.sgd_lr = None
        self.weight = _init_weight(self.K, out_features, in_features, dev, seed * 1000 + mp_comm.Get_rank())
        self._ops = ops
        self._build_ops()
        self.num_sms = 0

    def _build_ops(self):
        self.ar = self._ops.GemmAllReduce(self.comm, self.M, self.N) if self.mp > 1 else None

    def set_tokens(self, tokens: int) -> None:
        """Rebuild the fused op for another token count (COLLECTIVE; ``tp_layers.set_tokens``)."""
        set_tokens(self, tokens)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _RowFn.apply(x.reshape(self.M, -1), self.weight, self)
