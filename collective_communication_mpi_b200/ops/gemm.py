"""bf16 GEMMs on the hand-written tcgen05 kernel (csrc/gemm_tcgen05.cu).

The reference contains no matrix multiply: README.md:173-175 says forward/backward compute is
"taken care of", and func_impl.py only moves activations around it. These three products are
what a tensor-parallel FC layer needs between those communication steps, each in the storage
layout it naturally has (no transposes are ever materialised):

================  ==========================  =============================
``matmul_nn``     ``y  = x @ w``              forward,  w stored ``[in, out]``
``matmul_nt``     ``dx = dy @ w.T``           backward, w stored ``[in, out]``
``matmul_tn``     ``dw = x.T @ dy``           weight gradient
================  ==========================  =============================
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native as N


def _stream_ptr(stream=None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def _check_operand(t: torch.Tensor, name: str):
    if not t.is_cuda or t.dtype != torch.bfloat16 or t.dim() != 2:
        raise TypeError(f"{name} must be a 2-D CUDA bfloat16 tensor (got {t.dtype}, {tuple(t.shape)}, {t.device})")
    if t.stride(1) != 1:
        raise ValueError(f"{name} must have unit stride on its last axis (strides {t.stride()})")
    if t.stride(0) % 8 or t.data_ptr() % 16:
        raise ValueError(f"{name} needs 16-byte aligned rows (row stride {t.stride(0)} elements)")


EPI_NONE, EPI_GELU_DUAL, EPI_MUL_DGELU, EPI_ADD = range(4)      # EpiOp in csrc/gemm_tcgen05.cuh


def epilogue_args(M: int, Nn: int, device, *, gelu_out=None, dgelu_of=None, add=None):
    """Translate the three fused-epilogue keywords into ``(op, in_tensor, out2_tensor, ld)``.

    ``gelu_out``  a ``[M, N]`` bf16 tensor (or ``True`` to allocate one): the GEMM writes its result AND
                  ``gelu(result)`` there -- forward through an activation without a pass over HBM
    ``dgelu_of``  the saved pre-activation ``u``: result = ``(a @ b) * gelu'(u)`` -- backward through it
    ``add``       a ``[M, N]`` bf16 tensor: result = ``a @ b + add`` -- residual / gradient accumulation
    """
    given = [(EPI_GELU_DUAL, gelu_out), (EPI_MUL_DGELU, dgelu_of), (EPI_ADD, add)]
    given = [(op, t) for op, t in given if t is not None and t is not False]
    if not given:
        return EPI_NONE, None, None, 0
    if len(given) > 1:
        raise ValueError("at most one of gelu_out / dgelu_of / add")
    op, t = given[0]
    if t is True:
        t = torch.empty((M, Nn), dtype=torch.bfloat16, device=device)
    if t.dtype != torch.bfloat16 or t.dim() != 2 or tuple(t.shape) != (M, Nn) or t.stride(1) != 1 or t.stride(0) % 8:
        raise ValueError(f"epilogue operand must be a bfloat16 [{M}, {Nn}] tensor with unit column stride and 16-byte rows")
    if Nn % 32:
        raise ValueError("fused epilogues need N % 32 == 0")
    return (op, None, t, t.stride(0)) if op == EPI_GELU_DUAL else (op, t, None, t.stride(0))


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_km: bool = False, b_kn: bool = False,
         out: torch.Tensor | None = None, out_dtype: torch.dtype = torch.bfloat16,
         accumulate: bool = False, num_sms: int = 0, stream=None, gelu_out=None, dgelu_of=None, add=None,
         trace: torch.Tensor | None = None):
    """``C[M,N] = A[M,K] @ B[K,N]`` with fp32 accumulation in tensor memory.

    ``a`` is stored ``[M,K]`` (or ``[K,M]`` when ``a_km``); ``b`` is stored ``[N,K]`` (the
    ``nn.Linear`` weight layout) or ``[K,N]`` when ``b_kn``. ``num_sms`` caps the persistent grid
    (leave SMs free for a communication kernel running alongside). ``gelu_out`` / ``dgelu_of`` /
    ``add`` fuse elementwise work into the epilogue (``epilogue_args``); with ``gelu_out`` the call
    returns ``(C, gelu(C))``. ``trace``: a ``utils.ktrace`` buffer the kernel fills.
    """
    _check_operand(a, "a")
    _check_operand(b, "b")
    if a_km:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_kn:
        Kb, Nn = b.shape
    else:
        Nn, Kb = b.shape
    if K != Kb:
        raise ValueError(f"inner dimensions differ: {K} vs {Kb}")
    if out is None:
        out = torch.empty((M, Nn), dtype=out_dtype, device=a.device)
    else:
        if out.shape != (M, Nn) or out.stride(1) != 1 or not out.is_cuda:
            raise ValueError("out has the wrong shape / layout")
        out_dtype = out.dtype
    if out_dtype not in (torch.bfloat16, torch.float32):
        raise TypeError("GEMM output must be bfloat16 or float32")
    if accumulate and out_dtype != torch.float32:
        raise TypeError("accumulate=True needs a float32 output")
    op, e_in, e_out2, e_ld = epilogue_args(M, Nn, a.device, gelu_out=gelu_out, dgelu_of=dgelu_of, add=add)
    lib = N.device_lib()
    rc = lib.ccmpi_gemm_bf16_ex(
        C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(out.data_ptr()),
        M, Nn, K, a.stride(0), b.stride(0), out.stride(0), int(a_km), int(b_kn),
        N.BF16 if out_dtype == torch.bfloat16 else N.F32, int(accumulate), op,
        C.c_void_p(e_in.data_ptr() if e_in is not None else None),
        C.c_void_p(e_out2.data_ptr() if e_out2 is not None else None), e_ld,
        C.c_void_p(trace.data_ptr() if trace is not None else None), int(num_sms),
        C.c_void_p(_stream_ptr(stream)))
    N.check(rc, "ccmpi_gemm_bf16_ex")
    N.count_launch()
    return (out, e_out2) if op == EPI_GELU_DUAL else out


def matmul_nn(x: torch.Tensor, w: torch.Tensor, **kw) -> torch.Tensor:
    """``x[M,K] @ w[K,N]``."""
    return gemm(x, w, b_kn=True, **kw)


def matmul_nt(x: torch.Tensor, w: torch.Tensor, **kw) -> torch.Tensor:
    """``x[M,K] @ w[N,K].T``."""
    return gemm(x, w, **kw)


def matmul_tn(x: torch.Tensor, dy: torch.Tensor, **kw) -> torch.Tensor:
    """``x[T,M].T @ dy[T,N]`` (contraction over the token axis)."""
    return gemm(x, dy, a_km=True, b_kn=True, **kw)


def gemm_sgd_(w: torch.Tensor, a: torch.Tensor, b: torch.Tensor, lr: float, *, a_km: bool = False, b_kn: bool = False,
              step: int | None = None, seed: int = 0, num_sms: int = 0, stream=None) -> torch.Tensor:
    """``w -= lr * (a @ b)`` in ONE kernel: the weight-gradient GEMM applies the SGD step from its
    fp32 accumulator instead of writing the gradient (csrc/gemm_sgd.cu). Operand layouts as in
    ``gemm``; ``w`` is a bf16 ``[M, N]`` tensor with unit column stride, updated in place.
    With ``step`` (an optimizer step counter >= 0) the new weights are rounded stochastically
    (csrc/sr.cuh; bits from (row * ld + col, step, seed)); without it, to nearest."""
    _check_operand(a, "a")
    _check_operand(b, "b")
    K, M = a.shape if a_km else a.shape[::-1]
    Kb, Nn = b.shape if b_kn else b.shape[::-1]
    if K != Kb:
        raise ValueError(f"inner dimensions differ: {K} vs {Kb}")
    if w.dtype != torch.bfloat16 or w.shape != (M, Nn) or w.stride(1) != 1 or not w.is_cuda:
        raise ValueError("w must be a CUDA bfloat16 [M, N] tensor with unit column stride")
    rc = N.device_lib().ccmpi_gemm_sgd_bf16_sr(
        C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(w.data_ptr()), M, Nn, K, a.stride(0), b.stride(0),
        w.stride(0), int(a_km), int(b_kn), float(lr), -1 if step is None else int(step) & 0x7FFFFFFF, int(seed) & 0xFFFFFFFF,
        int(num_sms), C.c_void_p(_stream_ptr(stream)))
    N.check(rc, "ccmpi_gemm_sgd_bf16_sr")
    N.count_launch()
    return w
