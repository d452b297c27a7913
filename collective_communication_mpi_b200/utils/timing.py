"""Timing and bandwidth accounting.

The reference times with ``MPI.Wtime()`` around barrier-bracketed calls and prints a 100-run mean
(mpi-test.py:59-72,86-98). Here the rule is: device time (CUDA events on the launching stream),
after warm-up, synchronised on both sides, and the MAX over ranks for anything multi-GPU; host
paths use ``perf_counter``. Bandwidth uses the nccl-tests "bus bandwidth" convention so numbers are
comparable with NCCL's and with the 900 GB/s-per-direction NVLink figure.
"""
from __future__ import annotations

import time
from typing import Callable, Dict

BUS_FACTOR = {
    "allreduce": lambda n: 2.0 * (n - 1) / n,
    "allgather": lambda n: (n - 1) / n,
    "reduce_scatter": lambda n: (n - 1) / n,
    "alltoall": lambda n: (n - 1) / n,
}


def bus_bandwidth_gbs(op: str, nbytes: int, seconds: float, n: int) -> float:
    """GB/s on the busiest link for ``nbytes`` of payload (all-reduce: per-rank message; the
    others: total gathered / scattered bytes)."""
    return BUS_FACTOR[op](n) * nbytes / seconds / 1e9 if seconds > 0 else float("inf")


def time_device(fn: Callable[[], None], iters: int = 20, warmup: int = 3, comm=None, flush=None) -> float:
    """Milliseconds per call of ``fn`` on the current CUDA stream (max over ``comm``'s ranks).
    ``flush``: optional tensor larger than L2 that is overwritten between iterations."""
    import torch
    for _ in range(max(warmup, 1)):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    if flush is None:
        if comm is not None:
            comm.Barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        total = e0.elapsed_time(e1) / iters
    else:
        for _ in range(iters):
            flush.zero_()
            if comm is not None:
                comm.Barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            total += e0.elapsed_time(e1) / iters
    if comm is not None and comm.Get_size() > 1:
        total = max(comm.allgather(total))
    return total


def time_host(fn: Callable[[], None], iters: int = 100, warmup: int = 3, comm=None) -> float:
    """Seconds per call on the host, barrier-bracketed like mpi-test.py:59-64 (max over ranks)."""
    for _ in range(warmup):
        fn()
    acc = 0.0
    for _ in range(iters):
        if comm is not None:
            comm.Barrier()
        t0 = time.perf_counter()
        fn()
        if comm is not None:
            comm.Barrier()
        acc += time.perf_counter() - t0
    acc /= iters
    if comm is not None and comm.Get_size() > 1:
        acc = max(comm.allgather(acc))
    return acc


class OpStats:
    """Per-operation counters of a communicator: calls, payload bytes and (host-measured) seconds.
    Cheap enough to be always on; ``report()`` gives achieved GB/s per op."""

    def __init__(self):
        self.calls: Dict[str, int] = {}
        self.bytes: Dict[str, int] = {}
        self.seconds: Dict[str, float] = {}

    def add(self, op: str, nbytes: int, seconds: float) -> None:
        self.calls[op] = self.calls.get(op, 0) + 1
        self.bytes[op] = self.bytes.get(op, 0) + int(nbytes)
        self.seconds[op] = self.seconds.get(op, 0.0) + seconds

    def report(self) -> Dict[str, dict]:
        out = {}
        for op, c in self.calls.items():
            s = self.seconds[op]
            out[op] = {"calls": c, "bytes": self.bytes[op], "seconds": s,
                       "GBps": (self.bytes[op] / s / 1e9) if s > 0 else None}
        return out
