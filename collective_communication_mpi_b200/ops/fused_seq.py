"""Token-axis ("sequence parallel") forms of the fused ops: the activation is sharded by ROWS.

=====================  =====================================================  ====================
op                     computes                                               NCCL-style baseline
=====================  =====================================================  ====================
``ag_gemm_m``          ``concat_rows(x_0..x_{n-1}) @ w``                      all_gather ; matmul
``gemm_rs_m``          rows ``[r*M/n, (r+1)*M/n)`` of ``sum_s a_s @ b_s``     matmul ; reduce_scatter
=====================  =====================================================  ====================

The reference shards the feature axis only (model/func_impl.py:90,108,182); this is the layout
Megatron-style sequence parallelism adds on top. Kernels: csrc/fused_policies_m.cuh.

Run on B200s in round 2: emulated ranks on one GPU and real ranks on 2 / 4 / 8 (tests/test_gpu_seq.py);
numbers in profiles/r2_seq_bench_2gpu.jsonl. Used by ``parallel/sp_layers.py`` and, with
``CCMPI_FUSED_GATHER_OUTPUT=1``, by ``ColumnParallelLinear(gather_output=True)``.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from .. import _native as N
from .collectives import _arr, _sp
from .fused import _OUT, FusedOp, _ctx, _ptrs, ag_ctas


def ag_gemm_m_raw(comms, x_off: int, ldx: int, xfull_off: int, B: List[torch.Tensor], Cout: List[torch.Tensor],
                  M: int, Nn: int, K: int, *, b_kn: bool, flag_base: int, epoch: int, flag_target: int = 0,
                  num_sms: int = 0, stream=None):
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_ag_gemm_m(a, n, x_off, ldx, xfull_off, _ptrs(B), _ptrs(Cout), M, Nn, K,
                                              B[0].stride(0), Cout[0].stride(0), int(b_kn), _OUT[Cout[0].dtype],
                                              flag_base, epoch, flag_target & 0xFFFFFFFF, num_sms, _sp(stream))
    N.check(rc, "ccmpi_fused_ag_gemm_m")
    N.count_launch()


def gemm_rs_m_raw(comms, A: List[torch.Tensor], B: List[torch.Tensor], slots_off: int, Cout: List[torch.Tensor],
                  M: int, Nn: int, K: int, *, a_km: bool = False, b_kn: bool, flag_base: int, epoch: int,
                  num_sms: int = 0, stream=None):
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_gemm_rs_m(a, n, _ptrs(A), _ptrs(B), slots_off, _ptrs(Cout), M, Nn, K,
                                              A[0].stride(0), B[0].stride(0), Cout[0].stride(0), int(a_km), int(b_kn),
                                              _OUT[Cout[0].dtype], flag_base, epoch, num_sms, _sp(stream))
    N.check(rc, "ccmpi_fused_gemm_rs_m")
    N.count_launch()


def ag_m_flag_words(M: int) -> int:
    return M // 128


def rs_m_flag_words(M: int, Nfull: int, n: int) -> int:
    bn = 256 if Nfull % 256 == 0 else 128
    return (M // n // 128) * (Nfull // bn) * n


def _check_rows(M: int, n: int):
    if M % n or (M // n) % 256:
        raise ValueError(f"rows per rank M/n = {M}/{n} must be a multiple of 256")


class AllGatherGemmM(FusedOp):
    """``y = concat_rows(x_0, ..., x_{n-1}) @ w``; ``x``: this rank's ``[M/n, K]`` bf16 rows."""

    def __init__(self, comm, M: int, K: int):
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        _check_rows(M, n)
        self.M, self.K, self.rows = M, K, M // n
        self.x_off, self.x = h.empty((self.rows, K), torch.bfloat16)
        self.xfull = [h.empty((M, K), torch.bfloat16) for _ in range(2)]
        self.flag_base = self.ctx.new_slot(ag_m_flag_words(M))
        self.epoch = 0

    def input_buffer(self) -> torch.Tensor:
        return self.x

    def gathered(self) -> torch.Tensor:
        """The ``[M, K]`` activation the last call assembled on this rank (all rows, peers' included)."""
        return self.xfull[self.epoch & 1][1]

    def __call__(self, x: torch.Tensor, w: torch.Tensor, *, b_kn: bool = True, out: Optional[torch.Tensor] = None,
                 out_dtype=torch.bfloat16, num_sms: int = 0, stream=None) -> torch.Tensor:
        self._check_open()
        if x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x.reshape(self.rows, self.K))
        Nn = w.shape[1] if b_kn else w.shape[0]
        if out is None:
            out = torch.empty((self.M, Nn), dtype=out_dtype, device=w.device)
        self.epoch += 1
        self.flag_total = getattr(self, "flag_total", 0) + ag_ctas(1, self.M, Nn, 1, num_sms)
        ag_gemm_m_raw(self.ctx.dev.comms, self.x_off, self.K, self.xfull[self.epoch & 1][0], [w], [out], self.M, Nn,
                      self.K, b_kn=b_kn, flag_base=self.flag_base, epoch=self.epoch, flag_target=self.flag_total,
                      num_sms=num_sms, stream=stream)
        return out


class GemmReduceScatterM(FusedOp):
    """Rows ``[rank*M/n, (rank+1)*M/n)`` of ``sum over ranks of a @ b``; result ``[M/n, N]``."""

    def __init__(self, comm, M: int, Nfull: int, *, dtype=torch.bfloat16):
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        _check_rows(M, n)
        self.M, self.N, self.rows, self.dtype = M, Nfull, M // n, dtype
        self.slots = [h.empty((n, self.rows, Nfull), dtype) for _ in range(2)]
        self.flag_base = self.ctx.new_slot(rs_m_flag_words(M, Nfull, n))
        self.epoch = 0

    def __call__(self, a: torch.Tensor, b: torch.Tensor, *, a_km: bool = False, b_kn: bool = False,
                 out: Optional[torch.Tensor] = None, num_sms: int = 0, stream=None) -> torch.Tensor:
        self._check_open()
        K = a.shape[0] if a_km else a.shape[1]
        if out is None:
            out = torch.empty((self.rows, self.N), dtype=self.dtype, device=a.device)
        self.epoch += 1
        gemm_rs_m_raw(self.ctx.dev.comms, [a], [b], self.slots[self.epoch & 1][0], [out], self.M, self.N, K, a_km=a_km,
                      b_kn=b_kn, flag_base=self.flag_base, epoch=self.epoch, num_sms=num_sms, stream=stream)
        return out


# -------------------------------------------------------------------------------------------------
# gather -> GEMM -> gather (feature axis): the reference's complete forward step in one kernel
# -------------------------------------------------------------------------------------------------
def ag_gemm_ag_raw(comms, x_off: int, ldx: int, xfull_off: int, B: List[torch.Tensor], yfull_off: int, M: int, Nn: int,
                   K: int, *, b_kn: bool, out_dtype=torch.bfloat16, flag_base: int, epoch: int, flag_target: int = 0,
                   num_sms: int = 0, stream=None):
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_ag_gemm_ag(a, n, x_off, ldx, xfull_off, _ptrs(B), yfull_off, M, Nn, K,
                                               B[0].stride(0), int(b_kn), _OUT[out_dtype], flag_base, epoch,
                                               flag_target & 0xFFFFFFFF, num_sms, _sp(stream))
    N.check(rc, "ccmpi_fused_ag_gemm_ag")
    N.count_launch()


class AllGatherGemmAllGather(FusedOp):
    """``gather_last_axis(gather_last_axis(x) @ w)``: ``naive_collect_forward_input``, the FC GEMM
    and ``naive_collect_forward_output`` (func_impl.py:76-109) as ONE kernel. ``x``: ``[M, K/n]``
    bf16 per rank; ``w``: this rank's ``[K, N]`` (``b_kn``) or ``[N, K]`` shard; result ``[M, N*n]``
    in symmetric memory (valid until the call after next)."""

    def __init__(self, comm, M: int, K: int, Nlocal: int, *, dtype=torch.bfloat16):
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        if K % n or (K // n) % 64:
            raise ValueError(f"K/n = {K}/{n} must be a multiple of 64")
        self.M, self.K, self.N, self.chunk, self.dtype = M, K, Nlocal, K // n, dtype
        self.x_off, self.x = h.empty((M, self.chunk), torch.bfloat16)
        self.xfull = [h.empty((M, K), torch.bfloat16) for _ in range(2)]
        self.outs = [h.empty((M, Nlocal * n), dtype) for _ in range(2)]
        self.flag_base = self.ctx.new_slot(n * ((M + 127) // 128))
        self.epoch = 0

    def gathered_input(self) -> torch.Tensor:
        return self.xfull[self.epoch & 1][1]

    def __call__(self, x: torch.Tensor, w: torch.Tensor, *, b_kn: bool = True, num_sms: int = 0, stream=None) -> torch.Tensor:
        self._check_open()
        if x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x.reshape(self.M, self.chunk))
        self.epoch += 1
        self.flag_total = getattr(self, "flag_total", 0) + ag_ctas(0, self.M, self.N, 1, num_sms)
        off, out = self.outs[self.epoch & 1]
        ag_gemm_ag_raw(self.ctx.dev.comms, self.x_off, self.chunk, self.xfull[self.epoch & 1][0], [w], off, self.M,
                       self.N, self.K, b_kn=b_kn, out_dtype=self.dtype, flag_base=self.flag_base, epoch=self.epoch,
                       flag_target=self.flag_total, num_sms=num_sms, stream=stream)
        return out
