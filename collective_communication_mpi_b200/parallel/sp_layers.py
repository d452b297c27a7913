"""Tensor-parallel FC layers in the *sequence-parallel* activation layout (Megatron-SP).

Between the two GEMMs of a column/row pair the activation is feature-sharded exactly as in
``tp_layers.py``; OUTSIDE the pair it is sharded by tokens -- every rank holds ``M/n`` rows of the
full width -- so the elementwise work there (residuals, norms, dropout) shrinks by the group size:

``SeqColumnParallelLinear`` (weight ``[in, out/n]``)
    fwd  ``y = gather_rows(x_loc) @ W``              all-gather -> GEMM, gather axis = M
    bwd  ``dx_loc = rscatter_rows(dy @ W^T)``        GEMM -> reduce-scatter, scatter axis = M
         ``dW = gather_rows(x_loc)^T @ dy``          on the gathered rows kept from forward

``SeqRowParallelLinear`` (weight ``[in/n, out]``)
    fwd  ``y_loc = rscatter_rows(x @ W)``            GEMM -> reduce-scatter on M
    bwd  ``dx = gather_rows(dy_loc) @ W^T``          all-gather -> GEMM on M
         ``dW = x^T @ gather_rows(dy_loc)``

The reference never shards tokens inside a replica (its gathers and scatters are all on the feature
axis, func_impl.py:90,108,182); SURVEY section 5.7 lists this layout as the optional extension of the
same fused kernels. Kernels: ``ops/fused_seq.py``; without a GPU the layers
run the CPU statement in ``ops/host_ops.py``, which is how their logic is tested.
"""
from __future__ import annotations

import torch
from torch import nn

from ..ops import fused_seq as FS
from ..ops import host_ops as H
from .tp_layers import _backend, _dw_last, _init_weight, _weight_grad, set_tokens


def _seq_ops():
    return FS if torch.cuda.is_available() else H


class _SeqColumnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, layer):
        ctx.layer = layer
        y = layer.ag(x, weight, b_kn=True, num_sms=layer.num_sms)
        ctx.x_full = layer.ag.gathered()          # symmetric-heap view: kept as an attribute (see tp_layers)
        ctx.save_for_backward(weight)
        return y

    @staticmethod
    def backward(ctx, dy):
        (weight,) = ctx.saved_tensors
        layer, dy = ctx.layer, dy.contiguous()
        fused_sgd = _dw_last(layer)       # dW's kernel (or the reduction it triggers) rewrites W: dx first
        dw = None if fused_sgd else _weight_grad(layer, ctx.x_full, dy)
        dx = layer.rs(dy, weight, b_kn=False, num_sms=layer.num_sms)
        if fused_sgd:
            dw = _weight_grad(layer, ctx.x_full, dy)      # rewrites W: after dx has read it
        ctx.x_full = None
        return dx, dw, None


class _SeqRowFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, layer):
        ctx.layer = layer
        ctx.save_for_backward(x, weight)
        return layer.rs(x, weight, b_kn=True, num_sms=layer.num_sms)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        layer = ctx.layer
        dx = layer.ag(dy.contiguous(), weight, b_kn=False, num_sms=layer.num_sms)
        dw = _weight_grad(layer, x, layer.ag.gathered())
        return dx, dw, None


class SeqColumnParallelLinear(nn.Module):
    """``[M/n, in] -> [M, out/n]``: token-sharded in, feature-sharded out."""

    def __init__(self, mp_comm, tokens: int, in_features: int, out_features: int, *, grad_dtype=torch.bfloat16, seed: int = 0):
        super().__init__()
        self.comm, self.mp = mp_comm, mp_comm.Get_size()
        if self.mp < 2:
            raise ValueError("sequence-parallel layers need a model-parallel group of at least 2 (use tp_layers otherwise)")
        self.M, self.K, self.N = tokens, in_features, out_features // self.mp
        self.grad_dtype = grad_dtype
        dev, _, self._gemm, self._matmul_tn, self._gemm_sgd = _backend()
        self.sgd_lr = None
        self.weight = _init_weight(in_features, self.N, in_features, dev, seed * 1000 + mp_comm.Get_rank())
        self._build_ops()
        self.num_sms = 0

    def _build_ops(self):
        ops = _seq_ops()
        self.ag = ops.AllGatherGemmM(self.comm, self.M, self.K)
        self.rs = ops.GemmReduceScatterM(self.comm, self.M, self.K)

    def set_tokens(self, tokens: int) -> None:
        """Rebuild the fused ops for another token count (COLLECTIVE; ``tp_layers.set_tokens``)."""
        set_tokens(self, tokens)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _SeqColumnFn.apply(x.reshape(self.M // self.mp, -1), self.weight, self)


class SeqRowParallelLinear(nn.Module):
    """``[M, in/n] -> [M/n, out]``: feature-sharded in, token-sharded out (partial products reduced)."""

    def __init__(self, mp_comm, tokens: int, in_features: int, out_features: int, *, grad_dtype=torch.bfloat16, seed: int = 0):
        super().__init__()
        self.comm, self.mp = mp_comm, mp_comm.Get_size()
        if self.mp < 2:
            raise ValueError("sequence-parallel layers need a model-parallel group of at least 2 (use tp_layers otherwise)")
        self.M, self.K, self.N = tokens, in_features // self.mp, out_features
        self.grad_dtype = grad_dtype
        dev, _, self._gemm, self._matmul_tn, self._gemm_sgd = _backend()
        self.sgd_lr = None
        self.weight = _init_weight(self.K, out_features, in_features, dev, seed * 1000 + mp_comm.Get_rank())
        self._build_ops()
        self.num_sms = 0

    def _build_ops(self):
        ops = _seq_ops()
        self.rs = ops.GemmReduceScatterM(self.comm, self.M, self.N)
        self.ag = ops.AllGatherGemmM(self.comm, self.M, self.N)

    def set_tokens(self, tokens: int) -> None:
        """Rebuild the fused ops for another token count (COLLECTIVE; ``tp_layers.set_tokens``)."""
        set_tokens(self, tokens)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _SeqRowFn.apply(x.reshape(self.M, -1), self.weight, self)
