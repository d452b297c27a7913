"""The bar to beat (BASELINE.md B-NCCL-TP24): the same transformer-FC training step written the
way a PyTorch user would -- torch.distributed (NCCL) collectives + torch.matmul (cuBLAS), one op
after the other. Same layer math, same feature-sharded activation layout, same weights shapes,
same mixing / activation functions, same loss, same SGD as
collective_communication_mpi_b200.models.TrainStep; only the communication + GEMM kernels differ.

The gradient all-reduce is bucketed into one flat buffer and issued asynchronously on NCCL's
stream as soon as backward has produced it, which is what DDP does.
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
    def forward(ctx, x, w, group, mp):
        x_full = _gather_last(x, group, mp)
        ctx.save_for_backward(x_full, w)
        ctx.group, ctx.mp = group, mp
        return x_full @ w

    @staticmethod
    def backward(ctx, dy):
        x_full, w = ctx.saved_tensors
        dw = x_full.t() @ dy
        dx = _rscatter_last(dy @ w.t(), ctx.group, ctx.mp)
        return dx, dw, None, None


class _Row(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, group, mp):
        ctx.save_for_backward(x, w)
        ctx.group, ctx.mp = group, mp
        return _rscatter_last(x @ w, group, mp)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy_full = _gather_last(dy.contiguous(), ctx.group, ctx.mp)
        return dy_full @ w.t(), x.t() @ dy_full, None, None


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
        self.flat = torch.zeros(sum(p.numel() for p in self.params), dtype=torch.bfloat16, device=dev)
        self.views, off = [], 0
        for p in self.params:
            self.views.append(self.flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        # DDP-style overlap: a gradient is all-reduced (async, on NCCL's stream) as soon as
        # autograd has produced it, while backward continues with the earlier layers
        self.pending = []
        if dp > 1:
            for v, p in zip(self.views, self.params):
                p.register_post_accumulate_grad_hook(self._make_hook(v))

    def _make_hook(self, view):
        def hook(param):
            view.copy_(param.grad)
            self.pending.append(dist.all_reduce(view, group=self.dp_group, async_op=True))
        return hook

    def forward(self, x):
        q, k, v = _Column.apply(x, self.w_qkv, self.mp_group, self.mp).chunk(3, dim=-1)
        a = v * torch.sigmoid(q * k)
        h = x + _Row.apply(a, self.w_o, self.mp_group, self.mp)
        u = torch.nn.functional.gelu(_Column.apply(h, self.w_1, self.mp_group, self.mp), approximate="tanh")
        return h + _Row.apply(u, self.w_2, self.mp_group, self.mp)

    def __call__(self, x, target):
        for p in self.params:
            p.grad = None
        y = self.forward(x)
        loss = torch.nn.functional.mse_loss(y.float(), target.float())
        loss.backward()
        if self.dp > 1:
            for work in self.pending:
                work.wait()
            self.pending.clear()
            self.flat.mul_(1.0 / self.dp)
        else:
            for v, p in zip(self.views, self.params):
                v.copy_(p.grad)
        with torch.no_grad():
            torch._foreach_add_([p.data for p in self.params], self.views, alpha=-self.lr)
        return loss.detach()
