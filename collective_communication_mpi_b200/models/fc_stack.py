"""Model families built from the tensor-parallel FC layers.

``TransformerFCBlock`` -- the FC layers of one transformer block in the reference's naming
(fc_q / fc_k / fc_v / fc_o of ``get_info``, func_impl.py:65-70, plus fc1 / fc2 of BASELINE's
configs), feature-sharded activations in, feature-sharded activations out. The reference has no
attention (or any compute); the three projections are mixed by a cheap elementwise gate so every
weight receives a gradient.

``MLPStack`` -- a plain stack of square FC layers for the data-parallel configuration (1 B
parameters at 16 x 8192^2).

``TrainStep`` -- forward + backward + DP gradient reduction + SGD in one call.
"""
from __future__ import annotations

import os
from typing import List

import torch
from torch import nn

from ..ops import elementwise as EW
from ..parallel.dp import GradSink, sgd_step
from ..parallel.tp_layers import ColumnParallelLinear, RowParallelLinear, attn_block, mlp_block, set_tokens


class _Resizable:
    """``set_tokens(tokens)``: rebuild every layer's fused ops for another token count (COLLECTIVE over
    the model-parallel group; ``parallel.tp_layers.set_tokens``). The next ``forward`` takes inputs of the
    new size; weights and optimizer state are kept."""

    def set_tokens(self, tokens: int) -> None:
        set_tokens(*([self.tp_layers()[0], tokens] + self.tp_layers()[1:]))


class TransformerFCBlock(_Resizable, nn.Module):
    def __init__(self, mp_comm, tokens: int, hidden: int, ffn_mult: int = 4, *, seed: int = 0, num_comm_ctas: int = 0,
                 share_scatter: bool = False):
        super().__init__()
        self.hidden, self.mp = hidden, mp_comm.Get_size()
        kw = dict(num_comm_ctas=num_comm_ctas, share_scatter=share_scatter)
        # fc_q, fc_k, fc_v share their input: one gather, one GEMM with the three shards side by side
        self.fc_qkv = ColumnParallelLinear(mp_comm, tokens, hidden, 3 * hidden, seed=seed + 1, **kw)
        self.fc_o = RowParallelLinear(mp_comm, tokens, hidden, hidden, seed=seed + 2, **kw)
        self.fc1 = ColumnParallelLinear(mp_comm, tokens, hidden, ffn_mult * hidden, seed=seed + 3, **kw)
        self.fc2 = RowParallelLinear(mp_comm, tokens, ffn_mult * hidden, hidden, seed=seed + 4, **kw)
        # one-pass gate / GELU kernels with hand-written backward (CCMPI_FUSED_ELEMENTWISE=0: eager op chains)
        self.fused_elementwise = EW.enabled()
        # ... and GELU / GELU' / the residual adds inside the GEMM epilogues (CCMPI_FUSED_EPILOGUE=0: separate kernels)
        self.fused_epilogue = self.fused_elementwise and os.environ.get("CCMPI_FUSED_EPILOGUE", "1") != "0"

    def tp_layers(self) -> List[nn.Module]:
        return [self.fc_qkv, self.fc_o, self.fc1, self.fc2]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.fused_epilogue:
            return mlp_block(self.fc1, self.fc2, attn_block(self.fc_qkv, self.fc_o, x))
        if self.fused_elementwise:
            h = x + self.fc_o(EW.gate(self.fc_qkv(x)))
            return h + self.fc2(EW.gelu(self.fc1(h)))
        q, k, v = self.fc_qkv(x).chunk(3, dim=-1)
        a = v * torch.sigmoid(q * k)            # stand-in for attention: elementwise, keeps sharding
        h = x + self.fc_o(a)
        u = torch.nn.functional.gelu(self.fc1(h), approximate="tanh")
        return h + self.fc2(u)


class TransformerFCStack(_Resizable, nn.Module):
    """``depth`` transformer FC blocks in sequence (the reference's layer set fc_q/k/v/o + fc1/fc2,
    func_impl.py:65-70, repeated). All four reduce-scatters of a block produce ``[tokens, hidden]``, so with
    ``share_scatter_ops`` (default) the whole stack uses ONE GEMM->reduce-scatter op and its flag region
    (``parallel.tp_layers._scatter_op``): at 8192 tokens x 4096 over two ranks a block then takes 512 flag words
    instead of 4608 and 26 blocks fit a communicator's signal pad instead of 3."""

    def __init__(self, mp_comm, tokens: int, hidden: int, depth: int, ffn_mult: int = 4, *, seed: int = 0,
                 share_scatter_ops: bool = True):
        super().__init__()
        self.hidden, self.mp, self.depth = hidden, mp_comm.Get_size(), depth
        self.blocks = nn.ModuleList([TransformerFCBlock(mp_comm, tokens, hidden, ffn_mult, seed=seed + 10 * i,
                                                        share_scatter=share_scatter_ops) for i in range(depth)])

    def tp_layers(self) -> List[nn.Module]:
        return [l for b in self.blocks for l in b.tp_layers()]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for b in self.blocks:
            x = b(x)
        return x


class TransformerFCBlockSP(_Resizable, nn.Module):
    """The same block in the sequence-parallel layout (``parallel/sp_layers.py``): input and output are
    this rank's ``tokens/mp`` rows of the full width, so the residual adds touch 1/mp of the data; between
    the GEMM pairs the activation is feature-sharded as in ``TransformerFCBlock``."""

    def __init__(self, mp_comm, tokens: int, hidden: int, ffn_mult: int = 4, *, seed: int = 0):
        super().__init__()
        from ..parallel.sp_layers import SeqColumnParallelLinear, SeqRowParallelLinear
        self.hidden, self.mp = hidden, mp_comm.Get_size()
        self.fc_qkv = SeqColumnParallelLinear(mp_comm, tokens, hidden, 3 * hidden, seed=seed + 1)
        self.fc_o = SeqRowParallelLinear(mp_comm, tokens, hidden, hidden, seed=seed + 2)
        self.fc1 = SeqColumnParallelLinear(mp_comm, tokens, hidden, ffn_mult * hidden, seed=seed + 3)
        self.fc2 = SeqRowParallelLinear(mp_comm, tokens, ffn_mult * hidden, hidden, seed=seed + 4)
        self.fused_elementwise = EW.enabled()

    def tp_layers(self) -> List[nn.Module]:
        return [self.fc_qkv, self.fc_o, self.fc1, self.fc2]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.fused_elementwise:
            h = x + self.fc_o(EW.gate(self.fc_qkv(x)))
            return h + self.fc2(EW.gelu(self.fc1(h)))
        q, k, v = self.fc_qkv(x).chunk(3, dim=-1)
        h = x + self.fc_o(v * torch.sigmoid(q * k))
        u = torch.nn.functional.gelu(self.fc1(h), approximate="tanh")
        return h + self.fc2(u)


class TransformerFCBlockMegatron(_Resizable, nn.Module):
    """The same block in the classic Megatron layout (``parallel/megatron_layers.py``): the residual stream
    ``[tokens, hidden]`` is replicated on every model-parallel rank; each column/row pair ends in a GEMM fused
    with an all-reduce. Note that with replicated outputs every rank computes the same loss on the same data."""

    def __init__(self, mp_comm, tokens: int, hidden: int, ffn_mult: int = 4, *, seed: int = 0):
        super().__init__()
        from ..parallel.megatron_layers import MegatronColumnParallelLinear, MegatronRowParallelLinear
        self.hidden, self.mp = hidden, mp_comm.Get_size()
        self.fc_qkv = MegatronColumnParallelLinear(mp_comm, tokens, hidden, 3 * hidden, seed=seed + 1)
        self.fc_o = MegatronRowParallelLinear(mp_comm, tokens, hidden, hidden, seed=seed + 2)
        self.fc1 = MegatronColumnParallelLinear(mp_comm, tokens, hidden, ffn_mult * hidden, seed=seed + 3)
        self.fc2 = MegatronRowParallelLinear(mp_comm, tokens, ffn_mult * hidden, hidden, seed=seed + 4)
        self.fused_elementwise = EW.enabled()

    def tp_layers(self) -> List[nn.Module]:
        return [self.fc_qkv, self.fc_o, self.fc1, self.fc2]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.fused_elementwise:
            h = x + self.fc_o(EW.gate(self.fc_qkv(x)))
            return h + self.fc2(EW.gelu(self.fc1(h)))
        q, k, v = self.fc_qkv(x).chunk(3, dim=-1)
        h = x + self.fc_o(v * torch.sigmoid(q * k))
        u = torch.nn.functional.gelu(self.fc1(h), approximate="tanh")
        return h + self.fc2(u)


class MLPStack(_Resizable, nn.Module):
    def __init__(self, mp_comm, tokens: int, width: int, depth: int, *, seed: int = 0):
        super().__init__()
        self.layers = nn.ModuleList(
            [ColumnParallelLinear(mp_comm, tokens, width, width, seed=seed + i) for i in range(depth)])

    def tp_layers(self) -> List[nn.Module]:
        return list(self.layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for i, layer in enumerate(self.layers):
            x = layer(x)
            if i + 1 < len(self.layers):
                x = torch.relu(x)
        return x


class TrainStep:
    """One optimisation step of ``model`` on a 2-D grid: fused-TP forward/backward on ``mp_comm``
    (inside the layers), bucketed + overlapped gradient all-reduce on ``dp_comm``, SGD."""

    def __init__(self, model, dp_comm, *, lr: float = 1e-3, bucket_mb: int = 64, overlap: bool = True):
        self.model, self.lr = model, lr
        # without data parallelism every dW GEMM applies its own SGD step in the epilogue: no gradient
        # buffer, no reduction, no optimizer pass (CCMPI_FUSED_SGD=0 keeps them separate)
        self.fused_sgd = dp_comm.Get_size() == 1 and os.environ.get("CCMPI_FUSED_SGD", "1") != "0"
        if self.fused_sgd:
            for layer in model.tp_layers():
                layer.sgd_lr = lr
            self.sink = None
        else:
            self.sink = GradSink(dp_comm, model.tp_layers(), bucket_mb=bucket_mb, overlap=overlap)
        self.step_no = 0
        self.fused_loss = EW.enabled()

    def __call__(self, x: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        # the step counter selects this step's stochastic-rounding stream (csrc/sr.cuh)
        if self.fused_sgd:
            for layer in self.model.tp_layers():
                layer.sgd_step = self.step_no
        else:
            self.sink.begin_step(self.lr, self.step_no)
        self.step_no += 1
        y = self.model(x)
        if self.fused_loss and y.dtype == torch.bfloat16 and target.dtype == torch.bfloat16:
            loss, dy = EW.mse_loss_and_grad(y, target)      # loss and its gradient in one pass
            y.backward(dy)
        else:
            loss = torch.nn.functional.mse_loss(y.float(), target.float())
            loss.backward()
        if not self.fused_sgd:
            self.sink.apply()       # nothing left to do when the reductions were the optimizer step
        return loss.detach()
