"""Tensor-parallel FC layers on the fused kernels, in the reference's activation layout.

The reference keeps activations *feature-sharded*: every MP rank holds ``(b, s, d/mp)`` and the
communication steps gather / scatter the last axis (func_impl.py:76-187). With the weight
partitioning of ``get_info`` (func_impl.py:65-70) that gives two layer kinds, both mapping a
sharded input to a sharded output with exactly ONE fused kernel per GEMM:

``ColumnParallelLinear`` (fc_q / fc_k / fc_v / fc1; weight ``[in, out/mp]``)
    fwd  ``y_loc = gather(x_loc) @ W``          all-gather -> GEMM   (gather axis = K)
    bwd  ``dx_loc = rscatter(dy_loc @ W^T)``    GEMM -> reduce-scatter (scatter axis = N)
         ``dW = gather(x_loc)^T @ dy_loc``      plain GEMM on the gathered x kept from fwd

``RowParallelLinear`` (fc_o / fc2; weight ``[in/mp, out]``)
    fwd  ``y_loc = rscatter(x_loc @ W)``        GEMM -> reduce-scatter
    bwd  ``dx_loc = gather(dy_loc) @ W^T``      all-gather -> GEMM
         ``dW = x_loc^T @ gather(dy_loc)``      plain GEMM on the gathered dy kept from that kernel

(``naive_collect_forward_output`` -- gathering the *output* as well -- is available as
``gather_output=True``; a following layer would only re-shard it.)

Token count ``M = b*s`` is fixed per layer instance: staging buffers live in the symmetric heap
and are allocated once. Weights are plain ``[in, out]``-shaped bf16 arrays like the reference's.
All ranks must build their layers in the same order (symmetric allocation is collective by
convention).
"""
from __future__ import annotations

import math
import os

import torch
from torch import nn

from ..ops import fused as F
from ..ops import host_ops as H
from ..ops.gemm import gemm as _cuda_gemm, gemm_sgd_ as _cuda_gemm_sgd, matmul_tn as _cuda_matmul_tn


def _backend():
    """``(device, op classes, gemm, matmul_tn)``: the kernels whenever the process has a GPU, the
    CPU statement of the same ops (ops/host_ops.py) otherwise -- which is how the CPU test-suite
    runs these layers on real multi-process groups."""
    if torch.cuda.is_available():
        return torch.device("cuda", torch.cuda.current_device()), F, _cuda_gemm, _cuda_matmul_tn, _cuda_gemm_sgd
    return torch.device("cpu"), H, H.gemm, H.matmul_tn, H.gemm_sgd_


class _ColumnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, layer):
        ctx.layer = layer
        y, x_full = layer._fwd(x, weight)
        # x_full aliases the symmetric heap, whose views all share one autograd version counter
        # (any in-place write anywhere in the heap bumps it), so it is kept as a plain attribute;
        # its lifetime is managed by the layer's double buffering, not by autograd.
        ctx.x_full = x_full
        ctx.save_for_backward(weight)
        return y

    @staticmethod
    def backward(ctx, dy):
        (weight,) = ctx.saved_tensors
        # the first layer of a model gets data, not activations: nobody wants d loss / d input, so the
        # dx GEMM (and its reduce-scatter) is not run at all
        dx, dw = ctx.layer._bwd(dy.contiguous(), ctx.x_full, weight, need_dx=ctx.needs_input_grad[0])
        ctx.x_full = None
        return dx, dw, None


class _RowFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, layer):
        ctx.layer = layer
        ctx.save_for_backward(x, weight)
        return layer._fwd(x, weight)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        dx, dw = ctx.layer._bwd(dy.contiguous(), x, weight, need_dx=ctx.needs_input_grad[0])
        return dx, dw, None


def _weight_grad(layer, x, dy):
    """dW = x^T dy. With a gradient sink attached (``parallel/dp.py``) the GEMM writes straight
    into the data-parallel bucket in symmetric memory and the reducer is told the bucket entry is
    ready; autograd then gets ``None`` for the weight."""
    if layer.sgd_lr is not None:
        # no data-parallel reduction needed: the dW kernel applies the step itself (W = SR(W - lr * x^T dy))
        with torch.no_grad():
            layer._gemm_sgd(layer.weight.data, x, dy, layer.sgd_lr, a_km=True, b_kn=True, step=getattr(layer, "sgd_step", None),
                            seed=getattr(layer, "sgd_seed", 0), num_sms=layer.num_sms)
        return None
    sink = getattr(layer, "grad_sink", None)
    if sink is None:
        return layer._matmul_tn(x, dy, out_dtype=layer.grad_dtype, num_sms=layer.num_sms)
    layer._matmul_tn(x, dy, out=sink.view_for(layer.weight), num_sms=layer.num_sms)
    sink.ready(layer.weight)
    return None


def _dw_last(layer) -> bool:
    """dW must come AFTER dx when producing it lets a kernel rewrite W during backward: the dW GEMM with the
    SGD step in its epilogue, or a gradient sink whose reduction is the optimizer step. Otherwise dW goes
    first, so that its data-parallel bucket can start reducing while dx is computed."""
    if layer.sgd_lr is not None:
        return True
    sink = getattr(layer, "grad_sink", None)
    return sink is not None and sink.updates_in_backward


def set_tokens(layer, tokens: int, *more_layers) -> None:
    """Give a layer (or several: a whole block) another token count ``M``. The fused ops own staging
    buffers and flag regions sized by ``M`` (``[M, K]`` gather buffers, one flag per 128-row block), so
    the old ones are closed -- their symmetric memory and flag words go back to the allocators
    (``ops.fused.FusedOp.close``) -- and new ones are built in the constructors' order. With several
    layers everything is closed before anything is built, so growing does not need room for both
    generations. COLLECTIVE: every member of the layers' group calls it at the same point with the
    same ``tokens``. Weights, optimizer wiring and SM budgets are untouched."""
    layers = [l for l in (layer, *more_layers) if l.M != tokens]
    for l in layers:
        for name in ("ag", "rs", "ar", "out_ag"):
            op = getattr(l, name, None)
            if op is not None:
                op.close()
                setattr(l, name, None)
    for l in layers:
        l.M = tokens
        l._build_ops()


def _scatter_op(layer, width: int):
    """The layer's GEMM->reduce-scatter op for ``[M, width]`` results. With ``layer.share_scatter`` every
    layer of the communicator with that shape uses ONE op object: a reduce-scatter keeps nothing from one
    call to the next (slots are double-buffered by launch parity and a rank can never be two launches ahead
    of a peer whose partials it still reads, whichever layer launched), and its flag region -- one word per
    tile and source, 1024 words at 8192 x 4096 over two ranks -- is what runs out first in a deep stack
    (``N.PAD_USER_WORDS`` = 14336). The gather ops are NOT shared: their gathered operand is kept for the
    backward dW GEMM."""
    ops, comm = layer._ops, layer.comm
    if not getattr(layer, "share_scatter", False):
        return ops.GemmReduceScatter(comm, layer.M, width)
    cache = vars(comm).setdefault("_shared_scatter_ops", {})
    key = (ops.__name__, layer.M, width)
    op = cache.get(key)
    if op is None or op._closed:
        op = cache[key] = ops.GemmReduceScatter(comm, layer.M, width)
    else:
        op.retain()
    return op


def _init_weight(rows: int, cols: int, fan_in: int, device, seed: int) -> nn.Parameter:
    g = torch.Generator(device=device).manual_seed(seed)
    w = torch.randn(rows, cols, device=device, generator=g, dtype=torch.float32) * (1.0 / math.sqrt(fan_in))
    return nn.Parameter(w.to(torch.bfloat16))


class ColumnParallelLinear(nn.Module):
    """``[M, in/mp] -> [M, out/mp]`` (or ``[M, out]`` with ``gather_output``)."""

    def __init__(self, mp_comm, tokens: int, in_features: int, out_features: int, *, gather_output: bool = False,
                 grad_dtype=torch.bfloat16, seed: int = 0, num_comm_ctas: int = 0, share_scatter: bool = False):
        super().__init__()
        self.comm, self.mp = mp_comm, mp_comm.Get_size()
        self.M, self.K, self.N = tokens, in_features, out_features // self.mp
        self.gather_output, self.share_scatter = gather_output, share_scatter
        self.grad_dtype = grad_dtype
        dev, ops, self._gemm, self._matmul_tn, self._gemm_sgd = _backend()
        self.sgd_lr = None      # set by TrainStep when the optimizer step is fused into the dW GEMM
        self.sgd_step, self.sgd_seed = None, seed
        self.weight = _init_weight(in_features, self.N, in_features, dev, seed * 1000 + mp_comm.Get_rank())
        self._ops, self._dev, self._num_comm_ctas = ops, dev, num_comm_ctas
        self._build_ops()
        self.num_sms = 0
        self.bwd_num_sms = 0    # SM budget of the backward fused kernel (set by GradSink when a reduction runs next to it)

    def _build_ops(self):
        ops, comm, tokens, in_features = self._ops, self.comm, self.M, self.K
        # a 1-member group needs no staging buffers or flags: plain GEMMs
        self.ag = ops.AllGatherGemm(comm, tokens, in_features, num_comm_ctas=self._num_comm_ctas) if self.mp > 1 else None
        self.rs = _scatter_op(self, in_features) if self.mp > 1 else None
        # gather_output: by default the fused all-gather->GEMM is followed by the strided all-gather
        # kernel on its output; CCMPI_FUSED_GATHER_OUTPUT=1 selects the single gather->GEMM->gather
        # kernel (ops/fused_seq.py)
        self.out_ag = None
        if self.gather_output and self.mp > 1 and self._dev.type == "cuda" and os.environ.get("CCMPI_FUSED_GATHER_OUTPUT") == "1":
            from ..ops.fused_seq import AllGatherGemmAllGather
            self.out_ag = AllGatherGemmAllGather(comm, tokens, in_features, self.N)

    def set_tokens(self, tokens: int) -> None:
        """Rebuild the layer's fused ops for another token count (COLLECTIVE over the model-parallel
        group; see ``set_tokens`` below). The weight is untouched."""
        set_tokens(self, tokens)

    def input_buffer(self) -> torch.Tensor:
        """Write the layer input here (symmetric memory) to skip one device copy."""
        return self.ag.input_buffer() if self.ag is not None else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _ColumnFn.apply(x.reshape(self.M, -1), self.weight, self)

    # -- kernels ---------------------------------------------------------------------------
    def _fwd(self, x, w):
        if self.mp == 1:
            y = self._gemm(x, w, b_kn=True, num_sms=self.num_sms)
            return y, x
        if self.out_ag is not None:
            y = self.out_ag(x, w, b_kn=True, num_sms=self.num_sms)
            return y, self.out_ag.gathered_input()
        y = self.ag(x, w, b_kn=True, num_sms=self.num_sms)
        x_full = self.ag.gathered()
        if self.gather_output:
            # reference semantics (func_impl.py:94-109): also gather the output
            y = self.comm.gather_last_axis(y)
        return y, x_full

    def _bwd(self, dy, x_full, w, need_dx: bool = True):
        if self.gather_output and self.mp > 1:
            r = self.comm.Get_rank()
            dy = dy[:, r * self.N:(r + 1) * self.N].contiguous()     # func_impl.py:111-147
        # dW first: its data-parallel bucket can start reducing while dx is computed -- unless the
        # dW kernel itself rewrites W (fused SGD), in which case dx must read W before that
        late = _dw_last(self)
        dw = _weight_grad(self, x_full, dy) if not late else None
        if not need_dx:
            dx = None
        elif self.mp == 1:
            dx = self._gemm(dy, w, num_sms=self.num_sms)
        else:
            dx = self.rs(dy, w, b_kn=False, num_sms=self.bwd_num_sms)
        if late:
            dw = _weight_grad(self, x_full, dy)
        return dx, dw


class RowParallelLinear(nn.Module):
    """``[M, in/mp] -> [M, out/mp]``: partial products are reduce-scattered on the feature axis."""

    def __init__(self, mp_comm, tokens: int, in_features: int, out_features: int, *, grad_dtype=torch.bfloat16,
                 seed: int = 0, num_comm_ctas: int = 0, share_scatter: bool = False):
        super().__init__()
        self.comm, self.mp = mp_comm, mp_comm.Get_size()
        self.share_scatter = share_scatter
        self.M, self.K, self.N = tokens, in_features // self.mp, out_features
        self.grad_dtype = grad_dtype
        dev, ops, self._gemm, self._matmul_tn, self._gemm_sgd = _backend()
        self.sgd_lr = None
        self.sgd_step, self.sgd_seed = None, seed
        self.weight = _init_weight(self.K, out_features, in_features, dev, seed * 1000 + mp_comm.Get_rank())
        self._ops, self._num_comm_ctas = ops, num_comm_ctas
        self._build_ops()
        self.num_sms = 0
        self.bwd_num_sms = 0

    def _build_ops(self):
        ops, comm, tokens = self._ops, self.comm, self.M
        self.rs = _scatter_op(self, self.N) if self.mp > 1 else None
        self.ag = ops.AllGatherGemm(comm, tokens, self.N, num_comm_ctas=self._num_comm_ctas) if self.mp > 1 else None

    def set_tokens(self, tokens: int) -> None:
        """Rebuild the layer's fused ops for another token count (COLLECTIVE over the model-parallel group)."""
        set_tokens(self, tokens)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _RowFn.apply(x.reshape(self.M, -1), self.weight, self)

    def _fwd(self, x, w):
        if self.mp == 1:
            return self._gemm(x, w, b_kn=True, num_sms=self.num_sms)
        return self.rs(x, w, b_kn=True, num_sms=self.num_sms)

    def _bwd(self, dy, x, w, need_dx: bool = True):
        # (a row-parallel layer's dx kernel is also what gathers dy for the dW GEMM, so with mp > 1 it
        # runs whether or not dx is wanted)
        if self.mp == 1:
            dx = self._gemm(dy, w, num_sms=self.num_sms) if need_dx else None
            dy_full = dy
        else:
            dx = self.ag(dy, w, b_kn=False, num_sms=self.bwd_num_sms)
            dy_full = self.ag.gathered()
        dw = _weight_grad(self, x, dy_full)     # after dx: with fused SGD this kernel rewrites W
        return dx, dw


# -------------------------------------------------------------------------------------------------
# Two-layer blocks with the elementwise work between the GEMMs fused into their epilogues
# -------------------------------------------------------------------------------------------------
# A column-parallel layer followed by a row-parallel one is the unit a transformer block is made of
# (fc_q/k/v -> fc_o, fc1 -> fc2). Written as ONE autograd function per unit, the work between and
# around the two GEMMs rides in their epilogues instead of costing passes over HBM:
#
#   mlp_block   out = h + fc2(gelu(fc1(h)))
#       fwd  fc1 kernel writes u and gelu(u)            (EPI_GELU_DUAL)   -- no GELU kernel
#            fc2 kernel / its reducer adds h            (EPI_ADD)         -- no residual add
#       bwd  fc2's dx kernel multiplies by gelu'(u)     (EPI_MUL_DGELU)   -- no GELU-backward kernel
#            fc1's dx kernel / reducer adds d_out       (EPI_ADD)         -- no gradient-accumulation add
#   attn_block  h = x + fc_o(gate(fc_qkv(x)))           (gate = v * sigmoid(q * k), the stand-in for attention)
#       fwd  fc_o kernel / reducer adds x; bwd fc_qkv's dx kernel / reducer adds d_h
#
# Same kernels, same communication as the single layers (all-gather -> GEMM, GEMM -> reduce-scatter of
# func_impl.py:76-187); measured on a B200 the passes removed are ~0.47 ms of a 7.3 ms step.
def _col_fwd(layer, x, w, **epi):
    """Column-parallel forward ``gather(x) @ w`` -> ``(result or (u, act), x_full)``."""
    if layer.mp == 1:
        return layer._gemm(x, w, b_kn=True, num_sms=layer.num_sms, **epi), x
    y = layer.ag(x, w, b_kn=True, num_sms=layer.num_sms, **epi)
    return y, layer.ag.gathered()


def _col_dx(layer, dy, w, add):
    if layer.mp == 1:
        return layer._gemm(dy, w, num_sms=layer.num_sms, add=add)
    return layer.rs(dy, w, b_kn=False, num_sms=layer.bwd_num_sms, add=add)


def _row_fwd(layer, x, w, add):
    if layer.mp == 1:
        return layer._gemm(x, w, b_kn=True, num_sms=layer.num_sms, add=add)
    return layer.rs(x, w, b_kn=True, num_sms=layer.num_sms, add=add)


def _row_dx(layer, dy, w, **epi):
    """Row-parallel backward ``gather(dy) @ w^T`` -> ``(dx, dy_full)``."""
    if layer.mp == 1:
        return layer._gemm(dy, w, num_sms=layer.num_sms, **epi), dy
    dx = layer.ag(dy, w, b_kn=False, num_sms=layer.bwd_num_sms, **epi)
    return dx, layer.ag.gathered()


class _MlpBlockFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, w1, w2, fc1, fc2):
        h = h.reshape(fc1.M, -1)
        (u, act), h_full = _col_fwd(fc1, h, w1, gelu_out=True)
        out = _row_fwd(fc2, act, w2, add=h)
        ctx.fc1, ctx.fc2, ctx.h_full = fc1, fc2, h_full     # h_full may alias the symmetric heap: see _ColumnFn
        ctx.save_for_backward(u, act, w1, w2)
        return out

    @staticmethod
    def backward(ctx, dout):
        u, act, w1, w2 = ctx.saved_tensors
        fc1, fc2 = ctx.fc1, ctx.fc2
        dout = dout.contiguous()
        du, dout_full = _row_dx(fc2, dout, w2, dgelu_of=u)          # dx of fc2, through the activation
        dw2 = _weight_grad(fc2, act, dout_full)
        # dW before dx lets its data-parallel bucket start early -- unless the dW kernel rewrites W (fused SGD)
        late = _dw_last(fc1)
        dw1 = _weight_grad(fc1, ctx.h_full, du) if not late else None
        dh = _col_dx(fc1, du, w1, add=dout)                         # + the residual path's gradient
        if late:
            dw1 = _weight_grad(fc1, ctx.h_full, du)
        ctx.h_full = None
        return dh, dw1, dw2, None, None


class _AttnBlockFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, wqkv, wo, fc_qkv, fc_o):
        from ..ops import elementwise as EW
        x = x.reshape(fc_qkv.M, -1)
        qkv, x_full = _col_fwd(fc_qkv, x, wqkv)
        a = EW.gate_forward(qkv)
        h = _row_fwd(fc_o, a, wo, add=x)
        ctx.fc_qkv, ctx.fc_o, ctx.x_full = fc_qkv, fc_o, x_full
        ctx.save_for_backward(qkv, a, wqkv, wo)
        return h

    @staticmethod
    def backward(ctx, dh):
        from ..ops import elementwise as EW
        qkv, a, wqkv, wo = ctx.saved_tensors
        fc_qkv, fc_o = ctx.fc_qkv, ctx.fc_o
        dh = dh.contiguous()
        da, dh_full = _row_dx(fc_o, dh, wo)
        dwo = _weight_grad(fc_o, a, dh_full)
        dqkv = EW.gate_backward(da, qkv)
        need_dx = ctx.needs_input_grad[0]
        late = _dw_last(fc_qkv)
        dwqkv = _weight_grad(fc_qkv, ctx.x_full, dqkv) if not late else None
        dx = _col_dx(fc_qkv, dqkv, wqkv, add=dh) if need_dx else None
        if late:
            dwqkv = _weight_grad(fc_qkv, ctx.x_full, dqkv)
        ctx.x_full = None
        return dx, dwqkv, dwo, None, None


def _pair_ok(col, row):
    return isinstance(col, ColumnParallelLinear) and isinstance(row, RowParallelLinear) and not col.gather_output \
        and col.out_ag is None and col.N % 32 == 0 and row.N % 32 == 0


def mlp_block(fc1: "ColumnParallelLinear", fc2: "RowParallelLinear", h: torch.Tensor) -> torch.Tensor:
    """``h + fc2(gelu(fc1(h)))`` with GELU, its backward and both residual adds fused into the GEMMs."""
    if not _pair_ok(fc1, fc2):
        raise ValueError("mlp_block takes a ColumnParallelLinear (no gather_output) followed by a RowParallelLinear")
    return _MlpBlockFn.apply(h, fc1.weight, fc2.weight, fc1, fc2)


def attn_block(fc_qkv: "ColumnParallelLinear", fc_o: "RowParallelLinear", x: torch.Tensor) -> torch.Tensor:
    """``x + fc_o(gate(fc_qkv(x)))`` with the residual adds of both directions fused into the GEMMs."""
    if not _pair_ok(fc_qkv, fc_o):
        raise ValueError("attn_block takes a ColumnParallelLinear (no gather_output) followed by a RowParallelLinear")
    return _AttnBlockFn.apply(x, fc_qkv.weight, fc_o.weight, fc_qkv, fc_o)
