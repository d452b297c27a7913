"""The bar to beat (BASELINE.md B-NCCL-TP24): the same transformer-FC training step written the
way a PyTorch user would -- torch.distributed (NCCL) collectives + torch.matmul (cuBLAS), one op
after the other. Same layer math, same feature-sharded activation layout, same weights shapes,
same mixing / activation functions, same loss, same SGD as
collective_communication_mpi_b200.models.TrainStep; only the communication + GEMM kernels differ.

The gradient all-reduce is issued asynchronously on NCCL's stream as soon as backward has produced a
gradient, which is what DDP does. No avoidable work is charged to this arm: the dW GEMMs write straight
into their slice of the flat communication buffer (``torch.matmul(..., out=view)``: no packing copy, no
accumulation pass), the 1/dp average is folded into the learning rate (no scaling pass), and like our
arm it skips the input gradient of the first layer, which nobody consumes.
"""
from __future__ import annotations

import math

import torch
import torch.distributed as dist


def _gather_last(x, group, mp):
    if mp == 1:
        return x
    buf = torch.empty((mp,) + tuple(x.shape), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(buf, x.contiguous(), group=group)
    return buf.permute(1, 0, 2).reshape(x.shape[0], -1)


def _rscatter_last(y, group, mp):
    if mp == 1:
        return y
    m, n = y.shape
    blocks = y.view(m, mp, n // mp).permute(1, 0, 2).contiguous()
    out = torch.empty((m, n // mp), dtype=y.dtype, device=y.device)
    dist.reduce_scatter_tensor(out, blocks, group=group)
    return out


class _Column(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, group, mp, sink, idx):
        x_full = _gather_last(x, group, mp)
        ctx.save_for_backward(x_full, w)
        ctx.group, ctx.mp, ctx.sink, ctx.idx = group, mp, sink, idx
        return x_full @ w

    @staticmethod
    def backward(ctx, dy):
        x_full, w = ctx.saved_tensors
        torch.matmul(x_full.t(), dy, out=ctx.sink.views[ctx.idx])
        ctx.sink.grad_ready(ctx.idx)
        dx = _rscatter_last(dy @ w.t(), ctx.group, ctx.mp) if ctx.needs_input_grad[0] else None
        return dx, None, None, None, None, None


class _Row(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, group, mp, sink, idx):
        ctx.save_for_backward(x, w)
        ctx.group, ctx.mp, ctx.sink, ctx.idx = group, mp, sink, idx
        return _rscatter_last(x @ w, group, mp)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy_full = _gather_last(dy.contiguous(), ctx.group, ctx.mp)
        torch.matmul(x.t(), dy_full, out=ctx.sink.views[ctx.idx])
        ctx.sink.grad_ready(ctx.idx)
        dx = dy_full @ w.t() if ctx.needs_input_grad[0] else None
        return dx, None, None, None, None, None


class NcclTrainStep:
    def __init__(self, rank, world, mp, dp, tokens, hidden, ffn_mult, lr=1e-4):
        if not dist.is_initialized():
            import os
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29531")
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", torch.cuda.current_device()))
        self.mp, self.dp, self.lr = mp, dp, lr
        self.mp_group = self.dp_group = None
        for d in range(dp):
            g = dist.new_group(list(range(d * mp, (d + 1) * mp)))
            if rank // mp == d:
                self.mp_group = g
        for m in range(mp):
            g = dist.new_group(list(range(m, world, mp)))
            if rank % mp == m:
                self.dp_group = g
        dev = torch.device("cuda", torch.cuda.current_device())
        gen = torch.Generator(device=dev).manual_seed(rank)

        def w(rows, cols, fan_in):
            return (torch.randn(rows, cols, device=dev, generator=gen) / math.sqrt(fan_in)).bfloat16().requires_grad_(True)

        h, f = hidden, ffn_mult * hidden
        self.w_qkv, self.w_o = w(h, 3 * h // mp, h), w(h // mp, h, h)
        self.w_1, self.w_2 = w(h, f // mp, h), w(f // mp, h, f)
        self.params = [self.w_qkv, self.w_o, self.w_1, self.w_2]
        # the dW GEMMs write straight into their slice of one flat buffer, which is what NCCL reduces
        self.flat = torch.zeros(sum(p.numel() for p in self.params), dtype=torch.bfloat16, device=dev)
        self.views, off = [], 0
        for p in self.params:
            self.views.append(self.flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        self.pending = []

    def grad_ready(self, idx):
        """DDP-style overlap: a gradient is all-reduced (async, on NCCL's stream) as soon as backward has
        produced it, while backward continues with the earlier layers."""
        if self.dp > 1:
            self.pending.append(dist.all_reduce(self.views[idx], group=self.dp_group, async_op=True))

    def forward(self, x):
        g, mp = self.mp_group, self.mp
        q, k, v = _Column.apply(x, self.w_qkv, g, mp, self, 0).chunk(3, dim=-1)
        a = v * torch.sigmoid(q * k)
        h = x + _Row.apply(a, self.w_o, g, mp, self, 1)
        u = torch.nn.functional.gelu(_Column.apply(h, self.w_1, g, mp, self, 2), approximate="tanh")
        return h + _Row.apply(u, self.w_2, g, mp, self, 3)

    def __call__(self, x, target):
        y = self.forward(x)
        loss = torch.nn.functional.mse_loss(y.float(), target.float())
        loss.backward()
        for work in self.pending:
            work.wait()
        self.pending.clear()
        with torch.no_grad():       # SUM over replicas, averaged by the step size
            torch._foreach_add_([p.data for p in self.params], self.views, alpha=-self.lr / self.dp)
        return loss.detach()
