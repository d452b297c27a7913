"""Device-side trace of the persistent GEMM kernels: where each warp role of each CTA spent its time.

``ncu`` cannot look inside a kernel whose CTAs wait on another GPU (it replays one kernel on one
device), and a CUDA-event time says nothing about WHY a fused GEMM+collective kernel is slower than
the GEMM alone. With ``CCMPI_KERNEL_TRACE=1`` every fused op hands its kernel a small buffer
(16 x uint64 per CTA, ``TraceSlot`` in csrc/gemm_tcgen05.cuh); the elected thread of each role --
TMA producer, MMA issuer, epilogue -- adds up the SM cycles it spent blocked on each of its waits
and writes them out once at kernel end. ``summary()`` turns the last launch of an op into numbers
such as "producer waited 31 us for gathered chunks, the MMA pipe was starved for 28 us of 402 us".
Tracing costs a handful of predicated instructions per K-block when it is off.

The reference's only profiling is wall-clock ``MPI.Wtime`` around whole calls (mpi-test.py:59-72).
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

WORDS = 16
SLOTS = ["t0_ns", "prod_flag", "prod_empty", "mma_full", "mma_acc", "epi_bg", "epi_acc", "epi_tail", "t1_ns", "tiles"]
MAX_CTAS = 160 * 4
_buffers: Dict[int, torch.Tensor] = {}
_names: Dict[int, str] = {}


def enabled() -> bool:
    return os.environ.get("CCMPI_KERNEL_TRACE", "0") == "1"


def new_buffer(device=None) -> torch.Tensor:
    return torch.zeros(MAX_CTAS * WORDS, dtype=torch.int64, device=device or torch.device("cuda", torch.cuda.current_device()))


def buffer_for(owner, name: str) -> Optional[torch.Tensor]:
    """The trace buffer of op object ``owner`` (allocated on first use) when tracing is enabled, else None."""
    if not enabled():
        return None
    buf = _buffers.get(id(owner))
    if buf is None:
        buf = _buffers[id(owner)] = new_buffer()
        _names[id(owner)] = name
    buf.zero_()         # a launch with fewer CTAs must not leave the previous launch's rows behind
    return buf


def forget(owner) -> None:
    """Drop the trace buffer of an op that is being closed (``id()`` values are reused by later objects)."""
    _buffers.pop(id(owner), None)
    _names.pop(id(owner), None)


def last(owner) -> Optional[torch.Tensor]:
    """The buffer the last traced launch of ``owner`` wrote (no reset)."""
    return _buffers.get(id(owner))


def summary(buf: torch.Tensor, sm_mhz: float = 1965.0) -> dict:
    """Decode the last launch recorded in ``buf``: per-role blocked time, mean / max over CTAs, in us."""
    if buf.is_cuda:
        torch.cuda.synchronize()
    t = buf.view(-1, WORDS).cpu()
    live = t[:, 0] != 0
    t = t[live]
    if t.numel() == 0:
        return {"ctas": 0}
    cyc = 1.0 / sm_mhz          # us per cycle
    out = {"ctas": int(t.shape[0]), "kernel_us": round(float(t[:, 8].max() - t[:, 0].min()) / 1e3, 1),
           "start_skew_us": round(float(t[:, 0].max() - t[:, 0].min()) / 1e3, 1),
           "tiles_per_cta": round(float(t[:, 9].float().mean()), 2)}
    out["prod_slow_calls"] = round(float(t[:, 10].float().mean()), 1)
    out["prod_max_us"] = round(float(t[:, 11].float().max()) * cyc, 1)
    first = t[:, 12][t[:, 12] != 0]
    if first.numel():
        out["first_slow_wait_done_us"] = {"mean": round(float((first - t[:, 0].min()).float().mean()) / 1e3, 1),
                                          "max": round(float((first - t[:, 0].min()).float().max()) / 1e3, 1)}
    out["epi_slices"] = round(float(t[:, 13].float().mean()), 1)
    for i, name in enumerate(SLOTS):
        if name in ("t0_ns", "t1_ns", "tiles"):
            continue
        col = t[:, i].float() * cyc
        nz = col[col > 0]
        out[name + "_us"] = {"mean": round(float(nz.mean()), 1) if nz.numel() else 0.0,
                             "max": round(float(col.max()), 1)}
    return out


def report() -> dict:
    """``{op name #k: summary}`` for every op object that has traced a launch in this process."""
    res = {}
    for k, (oid, buf) in enumerate(_buffers.items()):
        res[f"{_names[oid]}#{k}"] = summary(buf)
    return res
