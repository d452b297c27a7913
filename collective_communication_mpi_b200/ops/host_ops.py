"""CPU statement of the GEMM and fused GEMM+collective ops: the same classes and call signatures as
``ops/gemm.py`` and ``ops/fused.py``, written as the unfused math -- library collective, then a
matmul in fp32 -- on CPU tensors over the shared-memory transport.

Two uses. (1) It is the reference's own dataflow (func_impl.py:76-187: gather, multiply, gather /
slice, multiply, reduce-scatter) in executable form, i.e. the oracle the CUDA kernels are compared
with. (2) The tensor-parallel layers, the models and the data-parallel gradient path pick these
implementations when the process has no GPU, so their Python logic -- autograd wiring, shard
bookkeeping, bucket scheduling -- is exercised by the CPU test-suite on real multi-process groups.

This is not a fallback for GPU boxes: with CUDA available the layers always use the kernels.
"""
from __future__ import annotations

from typing import Optional

import torch


def _epilogue(c: torch.Tensor, out_dtype, gelu_out, dgelu_of, add):
    """The fused epilogues of csrc/gemm_kernel.cuh on an fp32 product ``c``: returns ``(result, gelu or None)``."""
    from .elementwise import gelu_reference
    if dgelu_of is not None:
        from .elementwise import gelu_backward_reference
        c = gelu_backward_reference(c, dgelu_of.float()).float()
    if add is not None:
        c = c + add.float()
    y = c.to(out_dtype)
    if gelu_out is None or gelu_out is False:
        return y, None
    g = gelu_reference(y)           # from the rounded pre-activation, like the kernel
    if gelu_out is not True:
        gelu_out.copy_(g)
        g = gelu_out
    return y, g


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_km: bool = False, b_kn: bool = False, out: Optional[torch.Tensor] = None,
         out_dtype: torch.dtype = torch.bfloat16, accumulate: bool = False, num_sms: int = 0, stream=None,
         gelu_out=None, dgelu_of=None, add=None, trace=None):
    """Same contract as ``ops.gemm.gemm``: ``a`` is ``[M,K]`` (``[K,M]`` with ``a_km``), ``b`` is
    ``[N,K]`` (``[K,N]`` with ``b_kn``); fp32 accumulation, result cast to ``out_dtype``; the fused
    epilogue keywords are applied in fp32 before that cast."""
    A = a.float().t() if a_km else a.float()
    B = b.float() if b_kn else b.float().t()
    c = A @ B
    if out is not None:
        out_dtype = out.dtype
    y, g = _epilogue(c, out_dtype, gelu_out, dgelu_of, add)
    if out is not None:
        if accumulate:
            out.add_(y)
        else:
            out.copy_(y)
        y = out
    return (y, g) if g is not None else y


def matmul_tn(x: torch.Tensor, dy: torch.Tensor, **kw) -> torch.Tensor:
    return gemm(x, dy, a_km=True, b_kn=True, **kw)


def gemm_sgd_(w: torch.Tensor, a: torch.Tensor, b: torch.Tensor, lr: float, *, a_km: bool = False, b_kn: bool = False,
              step=None, seed: int = 0, num_sms: int = 0, stream=None) -> torch.Tensor:
    """``w -= lr * (a @ b)`` with the product kept in fp32 (what csrc/gemm_sgd.cu does in its epilogue); with
    ``step`` the result is rounded stochastically with the kernel's own arithmetic (utils/sr.py)."""
    new = w.float() - torch.tensor(lr, dtype=torch.float32) * gemm(a, b, a_km=a_km, b_kn=b_kn, out_dtype=torch.float32)
    if step is None:
        w.copy_(new.to(w.dtype))
    else:
        from ..utils import sr
        rows = torch.arange(w.shape[0], dtype=torch.int64).view(-1, 1) * w.stride(0)
        idx = rows + torch.arange(w.shape[1], dtype=torch.int64).view(1, -1)
        w.copy_(sr.round_bf16(new, idx, int(step) & 0x7FFFFFFF, seed))
    return w


class _HostOp:
    """Life cycle of the CUDA ops (``ops.fused.FusedOp``): the CPU statements own no symmetric memory or
    flag words, so ``close()`` only marks the object."""

    _closed = False
    _refs = 1

    def retain(self):
        self._refs += 1
        return self

    def close(self) -> None:
        if self._refs > 1:
            self._refs -= 1
            return
        self._closed = True

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


class AllGatherGemm(_HostOp):
    """``gather_last_axis(x) @ w`` (func_impl.py:76-91 + the FC GEMM)."""

    def __init__(self, comm, M: int, K: int, *, num_comm_ctas: int = 0, use_mc: bool = True):
        self.comm, self.M, self.K = comm, M, K
        self._gathered = None

    def gathered(self) -> torch.Tensor:
        return self._gathered

    def __call__(self, x, w, *, b_kn: bool = True, out=None, out_dtype=torch.bfloat16, num_sms: int = 0, stream=None,
                 gelu_out=None, dgelu_of=None, add=None):
        self._gathered = self.comm.gather_last_axis(x.reshape(self.M, -1).contiguous())
        return gemm(self._gathered, w, b_kn=b_kn, out=out, out_dtype=out_dtype, gelu_out=gelu_out, dgelu_of=dgelu_of, add=add)


class GemmReduceScatter(_HostOp):
    """``reduce_scatter_last_axis(a @ b)`` (the dx GEMM + func_impl.py:150-187)."""

    def __init__(self, comm, M: int, Nfull: int, *, dtype=torch.bfloat16):
        self.comm, self.M, self.N, self.dtype = comm, M, Nfull, dtype

    def __call__(self, a, b, *, a_km: bool = False, b_kn: bool = False, out=None, num_sms: int = 0, stream=None, add=None):
        part = gemm(a, b, a_km=a_km, b_kn=b_kn, out_dtype=torch.float32)
        red = self.comm.reduce_scatter_last_axis(part)
        if add is not None:
            red = red.float() + add.float()
        red = red.to(self.dtype)
        if out is not None:
            out.copy_(red)
            return out
        return red


class GemmAllReduce(_HostOp):
    """``all_reduce(a @ b)``."""

    def __init__(self, comm, M: int, Nfull: int, *, dtype=torch.bfloat16, use_mc: bool = True):
        self.comm, self.dtype = comm, dtype

    def __call__(self, a, b, *, a_km: bool = False, b_kn: bool = False, num_sms: int = 0, stream=None):
        part = gemm(a, b, a_km=a_km, b_kn=b_kn, out_dtype=torch.float32).contiguous()
        out = torch.empty_like(part)
        self.comm.Allreduce(part, out)
        return out.to(self.dtype)


class GemmAllGather(_HostOp):
    """``gather_last_axis(a @ b)`` (the FC GEMM + func_impl.py:94-109)."""

    def __init__(self, comm, M: int, Nlocal: int, *, dtype=torch.bfloat16, use_mc: bool = True):
        self.comm, self.dtype = comm, dtype

    def __call__(self, a, b, *, a_km: bool = False, b_kn: bool = True, num_sms: int = 0, stream=None):
        return self.comm.gather_last_axis(gemm(a, b, a_km=a_km, b_kn=b_kn, out_dtype=self.dtype))


# ---- token-axis (sequence-parallel) forms: ops/fused_seq.py -----------------------------------------
class AllGatherGemmM(_HostOp):
    """``concat_rows(x_0, ..., x_{n-1}) @ w``; ``x``: this rank's ``[M/n, K]`` rows."""

    def __init__(self, comm, M: int, K: int):
        self.comm, self.M, self.K = comm, M, K
        self._gathered = None

    def gathered(self) -> torch.Tensor:
        return self._gathered

    def __call__(self, x, w, *, b_kn: bool = True, out=None, out_dtype=torch.bfloat16, num_sms: int = 0, stream=None):
        full = torch.empty((self.M, self.K), dtype=x.dtype)
        self.comm.Allgather(x.reshape(-1, self.K).contiguous(), full)      # row blocks are contiguous: a flat all-gather
        self._gathered = full
        return gemm(full, w, b_kn=b_kn, out=out, out_dtype=out_dtype)


class GemmReduceScatterM(_HostOp):
    """Rows ``[rank*M/n, (rank+1)*M/n)`` of ``sum over ranks of a @ b``."""

    def __init__(self, comm, M: int, Nfull: int, *, dtype=torch.bfloat16):
        self.comm, self.M, self.N, self.dtype = comm, M, Nfull, dtype

    def __call__(self, a, b, *, a_km: bool = False, b_kn: bool = False, out=None, num_sms: int = 0, stream=None):
        part = gemm(a, b, a_km=a_km, b_kn=b_kn, out_dtype=torch.float32).contiguous()
        red = torch.empty((self.M // self.comm.Get_size(), self.N), dtype=torch.float32)
        self.comm.Reduce_scatter_block(part, red)
        red = red.to(self.dtype)
        if out is not None:
            out.copy_(red)
            return out
        return red
