"""Numerical self-check that ``bench.py`` runs in front of its timed region (every N; at N >= 2 it is
the multi-GPU correctness witness the timing lines cannot give).

One forward + backward of a small column-parallel -> row-parallel pair through the SAME code the
benchmark times -- fused all-gather->GEMM, GEMM->reduce-scatter, the dW GEMMs writing into the
data-parallel gradient buffer, the bucketed gradient all-reduce -- compared on every rank with an
oracle built from the framework's UNFUSED steps (``gather_last_axis`` / ``reduce_scatter_last_axis``,
i.e. the reference's ``naive_collect_forward_input`` and ``naive_collect_backward_x``,
model/func_impl.py:76-91,150-187) around fp32 ``torch.matmul``, and ``Allreduce`` over ``dp_comm``
(tests/test_get_info.py:39) for the gradients.

Returns ``{"max_rel_err": worst over ranks and tensors, "ok": bool, "what": {...}}``; relative error
is ``max|a - b| / max|b|`` per tensor. bf16 results of K <= 2048 sums: 2e-2 is several ulps of
headroom and far below what one missing partial (error ~ 1/mp) or a mis-scaled gradient would give.
"""
from __future__ import annotations

import torch

TOL = 2e-2


def run_selfcheck(world_comm, mp_comm, dp_comm, *, tokens: int = 1024, hidden: int = 1024) -> dict:
    from mpi4py import MPI
    from collective_communication_mpi_b200.parallel.dp import GradSink
    from collective_communication_mpi_b200.parallel.tp_layers import ColumnParallelLinear, RowParallelLinear

    mp, dp = mp_comm.Get_size(), dp_comm.Get_size()
    dev = torch.device("cuda", torch.cuda.current_device())
    col = ColumnParallelLinear(mp_comm, tokens, hidden, 2 * hidden, seed=71)
    row = RowParallelLinear(mp_comm, tokens, 2 * hidden, hidden, seed=72)
    sink = GradSink(dp_comm, [col, row])
    g = torch.Generator(device=dev).manual_seed(500 + world_comm.Get_rank())
    x = (torch.randn(tokens, hidden // mp, device=dev, generator=g) * 0.5).bfloat16().requires_grad_(True)
    dy = (torch.randn(tokens, hidden // mp, device=dev, generator=g) * 0.5).bfloat16()

    # ---- through the layers (fused kernels) ----
    # The weight gradients are checked through what the optimizer applied: (w_before - w_after) / lr. That covers
    # every data-parallel path alike -- including the one whose reduction kernel IS the optimizer step and never
    # materialises the averaged gradient. The step is large enough (>= 100 bf16 ulps of the weights) for the
    # stochastic rounding of the new weights to stay below the tolerance.
    LR = 0.01
    w_before = [col.weight.detach().float().clone(), row.weight.detach().float().clone()]
    sink.begin_step(LR, 0)
    h = col(x)
    y = row(h)
    y.backward(dy)
    sink.apply()
    torch.cuda.synchronize()
    got = {"y": y.detach().float(), "dx": x.grad.float(),
           "dW_col": (w_before[0] - col.weight.detach().float()) / LR, "dW_row": (w_before[1] - row.weight.detach().float()) / LR}

    # ---- oracle: unfused collectives + fp32 matmul ----
    with torch.no_grad():
        f32 = torch.float32
        wc, wr = w_before
        xf = mp_comm.gather_last_axis(x.detach()).to(f32) if mp > 1 else x.detach().to(f32)
        h_ref = (xf @ wc).bfloat16()                                   # the layer hands on a bf16 activation
        y_part = h_ref.to(f32) @ wr
        y_ref = mp_comm.reduce_scatter_last_axis(y_part.contiguous()) if mp > 1 else y_part
        dyf = mp_comm.gather_last_axis(dy).to(f32) if mp > 1 else dy.to(f32)
        dwr = h_ref.to(f32).t() @ dyf
        dh = (dyf @ wr.t()).bfloat16()
        dwc = xf.t() @ dh.to(f32)
        dx_part = dh.to(f32) @ wc.t()
        dx_ref = mp_comm.reduce_scatter_last_axis(dx_part.contiguous()) if mp > 1 else dx_part
        if dp > 1:
            for t in (dwr, dwc):
                tot = torch.empty_like(t)
                dp_comm.Allreduce(t.contiguous(), tot, op=MPI.SUM)
                t.copy_(tot / dp)
        ref = {"y": y_ref.float(), "dx": dx_ref.float(), "dW_col": dwc, "dW_row": dwr}
    errs = {k: float((got[k] - ref[k]).abs().max() / ref[k].abs().max().clamp_min(1e-12)) for k in ref}
    worst = world_comm.allgather(errs)
    merged = {k: max(e[k] for e in worst) for k in errs}
    m = max(merged.values())
    return {"max_rel_err": round(m, 6), "ok": bool(m < TOL), "tol": TOL, "per_tensor": {k: round(v, 6) for k, v in merged.items()},
            "shape": {"tokens": tokens, "hidden": hidden, "mp": mp, "dp": dp},
            "oracle": "unfused gather/reduce-scatter kernels + fp32 torch.matmul + Allreduce over dp_comm"}
