"""Host <-> device plumbing for a training loop: input prefetch and asynchronous result read-back.

The reference feeds NumPy arrays that already sit in host memory (``split_data`` returns views,
data_parallel_preprocess.py:52-57). On a GPU the equivalent is pinned host batches copied to the
device every step; done naively that copy (tens of MB per step here) and the per-step loss read
serialise with the kernels. These two helpers take both off the critical path without skipping
either: every step's inputs still cross PCIe and every step's loss still reaches the host.
"""
from __future__ import annotations

from typing import List, Sequence

import torch


class DevicePrefetcher:
    """Double-buffered host->device copies on a side stream.

    ``submit(host_tensors)`` starts copying one batch (pinned tensors); ``get()`` returns the device
    tensors of the oldest submitted batch once the compute stream may use them. A buffer is not
    overwritten before the kernels that read it have been enqueued and finished (tracked with
    events)."""

    def __init__(self, like: Sequence[torch.Tensor], depth: int = 2, device=None):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.depth = depth
        self.bufs: List[List[torch.Tensor]] = [[torch.empty(t.shape, dtype=t.dtype, device=self.device) for t in like]
                                              for _ in range(depth)]
        self.stream = torch.cuda.Stream(device=self.device)
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.free = [None] * depth
        self.head = self.tail = 0
        self.bytes_per_batch = sum(t.numel() * t.element_size() for t in like)

    def submit(self, host_tensors: Sequence[torch.Tensor]) -> None:
        i = self.head % self.depth
        self.head += 1
        with torch.cuda.stream(self.stream):
            if self.free[i] is not None:
                self.stream.wait_event(self.free[i])      # previous consumer of this buffer is done
            for dst, src in zip(self.bufs[i], host_tensors):
                dst.copy_(src, non_blocking=True)
            self.ready[i].record(self.stream)

    def get(self) -> List[torch.Tensor]:
        i = self.tail % self.depth
        self.tail += 1
        cur = torch.cuda.current_stream()
        cur.wait_event(self.ready[i])
        self._last = i
        return self.bufs[i]

    def release(self) -> None:
        """Call after the step that consumed the last ``get()`` has been enqueued."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.free[self._last] = ev


class AsyncScalarReader:
    """Device scalar -> pinned host memory without stalling the launch loop: ``push`` enqueues the
    copy, ``pop`` (normally one step later) waits for it and returns the Python float."""

    def __init__(self, depth: int = 4):
        self.host = torch.zeros(depth, dtype=torch.float32).pin_memory()
        self.events = [torch.cuda.Event() for _ in range(depth)]
        self.depth, self.head, self.tail = depth, 0, 0

    def push(self, value: torch.Tensor) -> None:
        i = self.head % self.depth
        self.head += 1
        self.host[i:i + 1].copy_(value.detach().reshape(1).float(), non_blocking=True)
        self.events[i].record(torch.cuda.current_stream())

    def pop(self) -> float:
        i = self.tail % self.depth
        self.tail += 1
        self.events[i].synchronize()
        return float(self.host[i])

    @property
    def pending(self) -> int:
        return self.head - self.tail
