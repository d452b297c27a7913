"""Fused GEMM + collective ops: ONE kernel per op computes tcgen05 tiles and moves operands /
results over NVLink from inside the kernel (csrc/fused_policies.cuh).

=====================  ==========================================================  ====================
op                     what it replaces in the reference's dataflow                NCCL-style baseline
=====================  ==========================================================  ====================
``ag_gemm``            ``naive_collect_forward_input`` + FC GEMM (func_impl:76)    all_gather ; matmul
``gemm_ag``            FC GEMM + ``naive_collect_forward_output`` (func_impl:94)   matmul ; all_gather
``gemm_rs``            dx GEMM + ``naive_collect_backward_x`` (func_impl:150)      matmul ; reduce_scatter
``gemm_ar``            row-parallel forward (fc_o / fc2): GEMM + all-reduce        matmul ; all_reduce
=====================  ==========================================================  ====================

``FusedContext`` owns what the kernels need between launches: the symmetric staging buffers
(double-buffered, because peers write into them), the flag region of the signal pad and the
launch epoch. One context per (communicator, op, shape); every rank must create and use its
contexts in the same order (allocation in the symmetric heap is collective by convention).

The ``*_raw`` functions take explicit device communicators and per-rank pointer lists; with more
than one communicator they run that many *emulated* ranks in one cooperative launch on one GPU,
which is how the single-GPU tests cover the cross-rank protocol.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from .. import _native as N
from ..utils import ktrace as KT
from .collectives import _arr, _sp
from .flags import FlagAllocator
from .gemm import EPI_GELU_DUAL, epilogue_args

_OUT = {torch.bfloat16: N.BF16, torch.float32: N.F32}


def _ptrs(tensors: Sequence[Optional[torch.Tensor]]):
    return (C.c_void_p * len(tensors))(*[(t.data_ptr() if t is not None else None) for t in tensors])


# -------------------------------------------------------------------------------------------------
# raw entry points
# -------------------------------------------------------------------------------------------------
def ag_gemm_raw(comms, x_off: int, ldx: int, xfull_off: int, B: List[torch.Tensor], Cout: List[torch.Tensor],
                M: int, Nn: int, K: int, *, b_kn: bool, flag_base: int, epoch: int, flag_target: int = 0,
                num_comm_ctas: int = 0, num_sms: int = 0, use_mc: bool = True, stream=None,
                epi_op: int = 0, epi_in: Optional[List[torch.Tensor]] = None,
                epi_out2: Optional[List[torch.Tensor]] = None, trace: Optional[torch.Tensor] = None):
    """``flag_target``: value the ready flags must reach for this launch (running sum of ``ag_ctas``
    over the launches on these flags); 0 = ``epoch`` times this launch's CTA count, which is right only
    when every launch on the flags has the same shape and SM budget (the tests' case)."""
    a, n = _arr(comms)
    e_t = epi_in if epi_in is not None else epi_out2
    rc = N.device_lib().ccmpi_fused_ag_gemm_ex(a, n, x_off, ldx, xfull_off, _ptrs(B), _ptrs(Cout), M, Nn, K,
                                               B[0].stride(0), Cout[0].stride(0), int(b_kn), _OUT[Cout[0].dtype],
                                               flag_base, epoch, flag_target & 0xFFFFFFFF, epi_op,
                                               _ptrs(epi_in) if epi_in is not None else None,
                                               _ptrs(epi_out2) if epi_out2 is not None else None,
                                               e_t[0].stride(0) if e_t is not None else 0,
                                               C.c_void_p(trace.data_ptr() if trace is not None else None),
                                               num_comm_ctas, num_sms, int(use_mc), _sp(stream))
    N.check(rc, "ccmpi_fused_ag_gemm_ex")
    N.count_launch()


def ag_ctas(kind: int, M: int, Nn: int, nvr: int, num_sms: int) -> int:
    """CTAs per rank of an all-gather->GEMM launch = contributions each of its ready flags gets from one
    source rank (kind 0: K-gather kernels, 1: M-gather)."""
    return N.check(N.device_lib().ccmpi_fused_ag_ctas(kind, M, Nn, nvr, num_sms), "ccmpi_fused_ag_ctas")


def gemm_rs_raw(comms, A: List[torch.Tensor], B: List[torch.Tensor], slots_off: int, Cout: List[torch.Tensor],
                M: int, Nn: int, K: int, *, a_km: bool = False, b_kn: bool, flag_base: int, epoch: int,
                num_sms: int = 0, stream=None, residual: Optional[List[torch.Tensor]] = None,
                trace: Optional[torch.Tensor] = None):
    """``residual``: one ``[M, Nn/n]`` bf16 tensor per rank, added to the reduced result by the reducer."""
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_gemm_rs_ex(a, n, _ptrs(A), _ptrs(B), slots_off, _ptrs(Cout), M, Nn, K,
                                               A[0].stride(0), B[0].stride(0), Cout[0].stride(0), int(a_km), int(b_kn),
                                               _OUT[Cout[0].dtype], flag_base, epoch,
                                               _ptrs(residual) if residual is not None else None,
                                               residual[0].stride(0) if residual is not None else 0,
                                               C.c_void_p(trace.data_ptr() if trace is not None else None),
                                               num_sms, _sp(stream))
    N.check(rc, "ccmpi_fused_gemm_rs_ex")
    N.count_launch()


def gemm_ar_raw(comms, A: List[torch.Tensor], B: List[torch.Tensor], slots_off: int, out_off: int, M: int, Nn: int,
                K: int, *, a_km: bool = False, b_kn: bool, out_dtype=torch.bfloat16, flag_base: int, epoch: int,
                num_sms: int = 0, use_mc: bool = True, stream=None):
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_gemm_ar(a, n, _ptrs(A), _ptrs(B), slots_off, out_off, M, Nn, K, A[0].stride(0),
                                            B[0].stride(0), int(a_km), int(b_kn), _OUT[out_dtype], flag_base, epoch,
                                            num_sms, int(use_mc), _sp(stream))
    N.check(rc, "ccmpi_fused_gemm_ar")
    N.count_launch()


def gemm_ag_raw(comms, A: List[torch.Tensor], B: List[torch.Tensor], yfull_off: int, M: int, Nn: int, K: int, *,
                a_km: bool = False, b_kn: bool, out_dtype=torch.bfloat16, num_sms: int = 0, use_mc: bool = True,
                stream=None):
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_gemm_ag(a, n, _ptrs(A), _ptrs(B), yfull_off, M, Nn, K, A[0].stride(0),
                                            B[0].stride(0), int(a_km), int(b_kn), _OUT[out_dtype], num_sms,
                                            int(use_mc), _sp(stream))
    N.check(rc, "ccmpi_fused_gemm_ag")
    N.count_launch()


# -------------------------------------------------------------------------------------------------
# stateful per-communicator contexts
# -------------------------------------------------------------------------------------------------
def _ag_flag_words(M: int, n: int) -> int:
    return n * ((M + 127) // 128)


def _rs_flag_words(M: int, chunk: int, n: int) -> int:
    bn = 256 if chunk % 256 == 0 else 128
    return ((M + 127) // 128) * (chunk // bn) * n


class FusedContext:
    """State shared by the fused ops on one communicator: flag-region and epoch bookkeeping."""

    def __init__(self, comm):
        self.comm = comm
        self.dev = comm.device_comm()
        self.heap = self.dev.heap
        self.n = comm.Get_size()
        self.rank = comm.Get_rank()
        self.flags = FlagAllocator(N.PAD_USER_WORDS)

    def new_slot(self, words: int) -> int:
        """Reserve ``words`` flag words of this communicator's signal pad. Flag regions of different
        live fused ops never overlap (each carries its own epochs); a region comes back through
        ``release_slot`` when its op is closed."""
        return self.flags.alloc(words)

    def release_slot(self, base: int) -> None:
        """Give a flag region back. COLLECTIVE over the communicator: every member closes the same
        op at the same point of its program.

        Flags are monotone ("wait until the word reaches this launch's target"), and what a region
        holds after its last launch depends on the op and its shape, so a new owner cannot simply
        carry on from there: the region is zeroed instead. Order: (1) this rank's launches that use
        the region have finished; (2) so have everybody's, hence no peer store into our words is
        still in flight; (3) zero our words; (4) nobody launches the next owner before every member
        has zeroed (a peer's first store must not be wiped by our late memset)."""
        words = self.flags.size_of(base)
        torch.cuda.synchronize()
        self.comm.Barrier()
        pad_off = self.dev._cid * N.PAD_BYTES + (N.PAD_USER + base) * 4
        self.heap.view(pad_off, (words,), torch.int32).zero_()
        torch.cuda.synchronize()
        self.comm.Barrier()
        self.flags.free(base)


class FusedOp:
    """Life cycle shared by the fused-op objects: symmetric buffers and a flag region are taken
    in ``__init__`` and given back by ``close()`` (collective; also a context manager). An op
    that is never closed keeps them for the life of the communicator, as before."""

    ctx: "FusedContext"
    flag_base: Optional[int] = None
    _closed = False
    _refs = 1

    def retain(self) -> "FusedOp":
        """Another holder of this op (layers of one shape may share an op whose buffers carry nothing from
        one call to the next, e.g. a GEMM->reduce-scatter: ``parallel.tp_layers`` with ``share_scatter``).
        Each holder calls ``close()`` once; the last one really closes."""
        self._refs += 1
        return self

    def _heap_offsets(self):
        """Offsets this op took from the symmetric heap: attributes named ``*_off`` and the
        ``[(offset, tensor), ...]`` lists ``heap.empty`` produced (double buffers)."""
        offs = set()
        for name, v in vars(self).items():
            if name.endswith("_off") and isinstance(v, int):
                offs.add(v)
            elif isinstance(v, list) and v and all(isinstance(i, tuple) and len(i) == 2 and isinstance(i[0], int)
                                                   and torch.is_tensor(i[1]) for i in v):
                offs.update(i[0] for i in v)
        return sorted(offs)

    def close(self) -> None:
        """Give the op's symmetric buffers and flag words back. COLLECTIVE: every member closes
        the same ops in the same order (``FusedContext.release_slot``)."""
        if self._closed:
            return
        if self._refs > 1:
            self._refs -= 1
            return
        self._closed = True
        if self.flag_base is not None:
            self.ctx.release_slot(self.flag_base)      # synchronises the device and the members
        else:
            torch.cuda.synchronize()
            self.ctx.comm.Barrier()
        for off in self._heap_offsets():
            self.ctx.heap.free(off)
        KT.forget(self)
        for name, v in list(vars(self).items()):       # drop the views: the memory may be handed out again
            if torch.is_tensor(v) or (isinstance(v, list) and v and isinstance(v[0], tuple)):
                setattr(self, name, None)

    def _check_open(self):
        if self._closed:
            raise RuntimeError(f"{type(self).__name__} used after close()")

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def close_fused_ops(module) -> int:
    """Close every fused op held by ``module`` and its submodules (COLLECTIVE; the order is the
    module tree's, which is the same on every member). Returns how many were closed. For programs
    that rebuild their layers with another token count: without this the old layers' flag words
    and symmetric buffers stay reserved until the communicator goes away."""
    done = 0
    for m in module.modules():
        for name in sorted(vars(m)):
            v = getattr(m, name)
            # the kernels' ops or their CPU statements (ops/host_ops.py): same life cycle
            if hasattr(v, "retain") and callable(getattr(v, "close", None)) and not getattr(v, "_closed", True):
                v.close()
                done += 1
    return done


def _ctx(comm) -> FusedContext:
    c = getattr(comm, "_fused_ctx", None)
    if c is None:
        c = FusedContext(comm)
        comm._fused_ctx = c
    return c


class AllGatherGemm(FusedOp):
    """``y = gather_last_axis(x) @ w`` with the gather fused into the GEMM's K loop.

    ``x``: ``[M, K/n]`` bf16 per rank (copied into / taken from the symmetric heap), ``w``:
    ``[K, N]`` (``b_kn=True``, the reference's weight layout) or ``[N, K]``."""

    def __init__(self, comm, M: int, K: int, *, num_comm_ctas: int = 0, use_mc: bool = True):
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        self.M, self.K, self.chunk = M, K, K // n
        self.x_off, self.x = h.empty((M, self.chunk), torch.bfloat16)
        self.xfull = [h.empty((M, K), torch.bfloat16) for _ in range(2)] if n > 1 else [h.empty((8, 8), torch.bfloat16)] * 2
        self.flag_base = self.ctx.new_slot(_ag_flag_words(M, n))
        self.epoch = 0
        self.flag_total = 0     # running sum of CTA counts: what the ready flags have reached after the last launch
        self.num_comm_ctas, self.use_mc = num_comm_ctas, use_mc and self.ctx.dev.has_mc

    def input_buffer(self) -> torch.Tensor:
        """Symmetric ``[M, K/n]`` tensor: producers that write their output here make the op
        zero-copy."""
        return self.x

    def gathered(self) -> torch.Tensor:
        """The ``[M, K]`` activation the last call assembled on this rank (valid until the call
        after next: the staging is double-buffered)."""
        return self.xfull[self.epoch & 1][1]

    def __call__(self, x: torch.Tensor, w: torch.Tensor, *, b_kn: bool = True, out: Optional[torch.Tensor] = None,
                 out_dtype=torch.bfloat16, num_sms: int = 0, stream=None, gelu_out=None, dgelu_of=None, add=None):
        """``gelu_out`` / ``dgelu_of`` / ``add``: elementwise work fused into the epilogue, as in
        ``ops.gemm.gemm``; with ``gelu_out`` the call returns ``(y, gelu(y))``."""
        self._check_open()
        if x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x.reshape(self.M, self.chunk))
        Nn = w.shape[1] if b_kn else w.shape[0]
        if out is None:
            out = torch.empty((self.M, Nn), dtype=out_dtype, device=w.device)
        op, e_in, e_out2, _ = epilogue_args(self.M, Nn, w.device, gelu_out=gelu_out, dgelu_of=dgelu_of, add=add)
        self.epoch += 1
        # every member launches the same shape with the same SM budget, so this rank's CTA count is
        # also the number of contributions its flags receive from each peer
        self.flag_total += ag_ctas(0, self.M, Nn, 1, num_sms)
        xf_off = self.xfull[self.epoch & 1][0]
        ag_gemm_raw(self.ctx.dev.comms, self.x_off, self.chunk, xf_off, [w], [out], self.M, Nn, self.K, b_kn=b_kn,
                    flag_base=self.flag_base, epoch=self.epoch, flag_target=self.flag_total,
                    num_comm_ctas=self.num_comm_ctas, num_sms=num_sms, use_mc=self.use_mc, stream=stream,
                    epi_op=op, epi_in=[e_in] if e_in is not None else None,
                    epi_out2=[e_out2] if e_out2 is not None else None, trace=KT.buffer_for(self, "ag_gemm"))
        return (out, e_out2) if op == EPI_GELU_DUAL else out


class GemmReduceScatter(FusedOp):
    """``reduce_scatter_last_axis(a @ b)`` with the scatter + ordered reduction fused into the
    GEMM epilogue / tail. Result ``[M, N/n]``."""

    def __init__(self, comm, M: int, Nfull: int, *, dtype=torch.bfloat16):
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        self.M, self.N, self.chunk, self.dtype = M, Nfull, Nfull // n, dtype
        self.slots = [h.empty((n, M, self.chunk), dtype) for _ in range(2)]
        self.flag_base = self.ctx.new_slot(_rs_flag_words(M, self.chunk, n))
        self.epoch = 0

    def __call__(self, a: torch.Tensor, b: torch.Tensor, *, a_km: bool = False, b_kn: bool = False,
                 out: Optional[torch.Tensor] = None, num_sms: int = 0, stream=None,
                 add: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``add``: a ``[M, N/n]`` bf16 tensor added to the reduced result by the reducer (residual
        connection / gradient accumulation) before the single rounding."""
        self._check_open()
        K = a.shape[0] if a_km else a.shape[1]
        if out is None:
            out = torch.empty((self.M, self.chunk), dtype=self.dtype, device=a.device)
        if add is not None and (add.dtype != torch.bfloat16 or tuple(add.shape) != (self.M, self.chunk) or add.stride(1) != 1):
            raise ValueError(f"add must be a bfloat16 [{self.M}, {self.chunk}] tensor with unit column stride")
        self.epoch += 1
        gemm_rs_raw(self.ctx.dev.comms, [a], [b], self.slots[self.epoch & 1][0], [out], self.M, self.N, K, a_km=a_km,
                    b_kn=b_kn, flag_base=self.flag_base, epoch=self.epoch, num_sms=num_sms, stream=stream,
                    residual=[add] if add is not None else None, trace=KT.buffer_for(self, "gemm_rs"))
        return out


class GemmAllReduce(FusedOp):
    """``all_reduce(a @ b)``: every rank ends with the full ``[M, N]`` sum in a symmetric buffer."""

    def __init__(self, comm, M: int, Nfull: int, *, dtype=torch.bfloat16, use_mc: bool = True):
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        self.M, self.N, self.chunk, self.dtype = M, Nfull, Nfull // n, dtype
        self.slots = [h.empty((n, M, self.chunk), dtype) for _ in range(2)]
        self.outs = [h.empty((M, Nfull), dtype) for _ in range(2)]
        self.flag_base = self.ctx.new_slot(_rs_flag_words(M, self.chunk, n))
        self.epoch = 0
        self.use_mc = use_mc and self.ctx.dev.has_mc

    def __call__(self, a: torch.Tensor, b: torch.Tensor, *, a_km: bool = False, b_kn: bool = False, num_sms: int = 0,
                 stream=None) -> torch.Tensor:
        self._check_open()
        K = a.shape[0] if a_km else a.shape[1]
        self.epoch += 1
        off, out = self.outs[self.epoch & 1]
        gemm_ar_raw(self.ctx.dev.comms, [a], [b], self.slots[self.epoch & 1][0], off, self.M, self.N, K, a_km=a_km,
                    b_kn=b_kn, out_dtype=self.dtype, flag_base=self.flag_base, epoch=self.epoch, num_sms=num_sms,
                    use_mc=self.use_mc, stream=stream)
        return out


class GemmAllGather(FusedOp):
    """``gather_last_axis(a @ b)``: each finished tile is stored into every rank's ``[M, N*n]``."""

    def __init__(self, comm, M: int, Nlocal: int, *, dtype=torch.bfloat16, use_mc: bool = True):
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        self.M, self.N, self.dtype = M, Nlocal, dtype
        self.outs = [h.empty((M, Nlocal * n), dtype) for _ in range(2)]
        self.epoch = 0
        self.use_mc = use_mc and self.ctx.dev.has_mc

    def __call__(self, a: torch.Tensor, b: torch.Tensor, *, a_km: bool = False, b_kn: bool = True, num_sms: int = 0,
                 stream=None) -> torch.Tensor:
        self._check_open()
        K = a.shape[0] if a_km else a.shape[1]
        self.epoch += 1
        off, out = self.outs[self.epoch & 1]
        gemm_ag_raw(self.ctx.dev.comms, [a], [b], off, self.M, self.N, K, a_km=a_km, b_kn=b_kn, out_dtype=self.dtype,
                    num_sms=num_sms, use_mc=self.use_mc, stream=stream)
        return out
