"""Thin, allocation-free wrappers over the collective kernels (csrc/collectives.cu).

Everything here works on *heap offsets*: the caller has already placed its operands in the
symmetric heap (see runtime/symm.py). ``comms`` is a list of device communicators -- one in
production, several when emulating ranks on a single GPU -- and is turned into the contiguous
C array the native entry points expect.

Reference call sites these replace: comm.py:16 (Barrier), :22 (Allreduce), :29 (Allgather),
:36 (Reduce_scatter_block), :61 (Alltoall), :63-107 (myAllreduce), :110-199 (myAlltoall/2).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

import torch

from .. import _native as N

_TORCH2CC = None


def cc_dtype(dtype) -> int:
    """torch / numpy dtype -> ccmpi_dtype."""
    global _TORCH2CC
    if _TORCH2CC is None:
        import numpy as np
        _TORCH2CC = {
            torch.float32: N.F32, torch.bfloat16: N.BF16, torch.float16: N.F16, torch.float64: N.F64,
            torch.int32: N.I32, torch.int64: N.I64, torch.uint8: N.U8,
            np.dtype("float32"): N.F32, np.dtype("float16"): N.F16, np.dtype("float64"): N.F64,
            np.dtype("int32"): N.I32, np.dtype("int64"): N.I64, np.dtype("uint8"): N.U8,
        }
    try:
        return _TORCH2CC[dtype]
    except (KeyError, TypeError):
        import numpy as np
        try:
            return _TORCH2CC[np.dtype(dtype)]
        except Exception:
            raise TypeError(f"dtype {dtype!r} is not supported by the device collectives") from None


def _arr(comms: Sequence[N.DevComm]):
    n = len(comms)
    if not 1 <= n <= N.MAX_VRANKS:
        raise ValueError(f"{n} emulated ranks requested; supported: 1..{N.MAX_VRANKS}")
    a = (N.DevComm * n)(*comms)
    return a, n


def _sp(stream) -> C.c_void_p:
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def barrier(comms, stream=None):
    a, n = _arr(comms)
    N.check(N.device_lib().ccmpi_barrier(a, n, _sp(stream)), "ccmpi_barrier")
    N.count_launch()


def allreduce(comms, dtype: int, op: int, src_off: int, dst_off: int, count: int, *, tmp_off: int = 0,
              algo: int = N.AR_AUTO, scale: float = 1.0, out_dtype: int = -1, max_blocks: int = 0, stream=None,
              tight=None):
    """``tight`` (benchmarks / tests): True forces the group-size-specialised two-shot kernel, False the
    generic one; None lets the library choose (the specialised kernel whenever it applies)."""
    a, n = _arr(comms)
    if tight is True and algo == N.AR_TWOSHOT:
        N.check(N.device_lib().ccmpi_allreduce_twoshot_tight(a, n, dtype, op, src_off, dst_off, tmp_off, count,
                                                             C.c_float(scale), max_blocks, _sp(stream)),
                "ccmpi_allreduce_twoshot_tight")
        N.count_launch()
        return
    if tight is False and algo == N.AR_TWOSHOT:
        algo = N.AR_TWOSHOT_GENERIC
    N.check(N.device_lib().ccmpi_allreduce(a, n, algo, dtype, op, src_off, dst_off, tmp_off, count,
                                           C.c_float(scale), out_dtype, max_blocks, _sp(stream)), "ccmpi_allreduce")
    N.count_launch()


def allreduce_sgd(comms, grad_off: int, w_off: int, count: int, lr_eff: float, *, base_index: int, step: int, seed: int = 0,
                  max_blocks: int = 0, stream=None):
    """Data-parallel gradient all-reduce fused with the SGD update (NVLS, bf16): see ``ccmpi_allreduce_sgd_nvls``."""
    a, n = _arr(comms)
    N.check(N.device_lib().ccmpi_allreduce_sgd_nvls(a, n, grad_off, w_off, count, C.c_float(lr_eff), int(base_index),
                                                    int(step) & 0xFFFFFFFF, int(seed) & 0xFFFFFFFF, max_blocks, _sp(stream)),
            "ccmpi_allreduce_sgd_nvls")
    N.count_launch()


def allgather(comms, src_off: int, dst_off: int, rows: int, row_bytes: int, src_row_stride: int,
              dst_row_stride: int, *, use_mc: bool = True, max_blocks: int = 0, stream=None):
    a, n = _arr(comms)
    # use_mc: False = per-peer stores, True = library's choice (multicast below ~8 MB), 2 = multicast at any size
    N.check(N.device_lib().ccmpi_allgather(a, n, src_off, dst_off, rows, row_bytes, src_row_stride,
                                           dst_row_stride, int(use_mc), max_blocks, _sp(stream)), "ccmpi_allgather")
    N.count_launch()


def reduce_scatter(comms, dtype: int, op: int, src_off: int, dst_off: int, rows: int, row_elems: int,
                   src_row_stride: int, dst_row_stride: int, *, use_mc: bool = True, max_blocks: int = 0,
                   stream=None, tight=None):
    """``tight`` as in ``allreduce``: True / False force the specialised / generic peer-load kernel."""
    a, n = _arr(comms)
    if tight is True:
        N.check(N.device_lib().ccmpi_reduce_scatter_tight(a, n, dtype, op, src_off, dst_off, rows, row_elems,
                                                          src_row_stride, dst_row_stride, max_blocks, _sp(stream)),
                "ccmpi_reduce_scatter_tight")
        N.count_launch()
        return
    mode = 2 if tight is False and not use_mc else int(use_mc)     # use_mc: False / True (library's choice) / 3 (NVLS at any size)
    N.check(N.device_lib().ccmpi_reduce_scatter(a, n, dtype, op, src_off, dst_off, rows, row_elems,
                                                src_row_stride, dst_row_stride, mode, max_blocks,
                                                _sp(stream)), "ccmpi_reduce_scatter")
    N.count_launch()


def alltoall(comms, src_off: int, dst_off: int, seg_bytes: int, *, staggered: bool = False, max_blocks: int = 0,
             stream=None):
    a, n = _arr(comms)
    N.check(N.device_lib().ccmpi_alltoall(a, n, src_off, dst_off, seg_bytes, int(staggered), max_blocks,
                                          _sp(stream)), "ccmpi_alltoall")
    N.count_launch()


def put_signal(comm: N.DevComm, peer: int, src_off: int, dst_off: int, nbytes: int, flag: int, stream=None):
    a, n = _arr([comm])
    N.check(N.device_lib().ccmpi_put_signal(a, n, peer, src_off, dst_off, nbytes, flag, _sp(stream)), "ccmpi_put_signal")
    N.count_launch()


def wait_signal(comm: N.DevComm, peer: int, flag: int, stream=None):
    a, n = _arr([comm])
    N.check(N.device_lib().ccmpi_wait_signal(a, n, peer, flag, _sp(stream)), "ccmpi_wait_signal")
    N.count_launch()
