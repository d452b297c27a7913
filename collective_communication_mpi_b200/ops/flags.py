"""Allocator for the flag words of a communicator's signal pad.

The fused GEMM+collective kernels signal each other through 32-bit words in the user area of the
signal pad (``CCMPI_PAD_USER`` onwards, ``csrc/fused_policies.cuh: user_flag``). Every fused-op
object owns a run of those words for its whole life; when a model is rebuilt with other shapes
(another sequence length, another grid) the old objects are closed and their runs come back here.

Pure Python on purpose (no torch, no CUDA): the policy is tested on the CPU
(``tests/test_flag_allocator.py``); what a release has to do on the device (quiesce, zero the
words on every member) lives in ``ops.fused.FusedContext.release_slot``.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

GRANULE = 16  # words; one 64-byte sector, so two ops never share a sector


class FlagSpaceExhausted(RuntimeError):
    pass


class FlagAllocator:
    """First-fit allocator over ``[0, capacity)`` words with coalescing of freed runs.

    With nothing freed it behaves exactly like a bump pointer, so the word layout of a program
    that never closes an op is the layout it always had."""

    def __init__(self, capacity: int):
        if capacity <= 0 or capacity % GRANULE:
            raise ValueError(f"capacity must be a positive multiple of {GRANULE} words")
        self.capacity = capacity
        self._free: List[Tuple[int, int]] = [(0, capacity)]
        self._live: Dict[int, int] = {}

    @staticmethod
    def rounded(words: int) -> int:
        return max(GRANULE, (int(words) + GRANULE - 1) // GRANULE * GRANULE)

    def alloc(self, words: int) -> int:
        size = self.rounded(words)
        for i, (base, run) in enumerate(self._free):
            if run >= size:
                if run == size:
                    self._free.pop(i)
                else:
                    self._free[i] = (base + size, run - size)
                self._live[base] = size
                return base
        raise FlagSpaceExhausted(
            f"signal pad exhausted: need {size} flag words, largest free run "
            f"{max((r for _, r in self._free), default=0)} of {self.capacity} "
            f"({self.used} in use by {len(self._live)} fused ops); close() fused ops that are no longer "
            "used, share them between layers of the same shape, or use fewer tokens per call")

    def size_of(self, base: int) -> int:
        return self._live[base]

    def free(self, base: int) -> int:
        """Return the run that starts at ``base``; returns its size in words."""
        try:
            size = self._live.pop(base)
        except KeyError:
            raise ValueError(f"flag run at word {base} is not allocated (double close?)") from None
        self._free.append((base, size))
        self._free.sort()
        merged: List[Tuple[int, int]] = []
        for b, s in self._free:
            if merged and merged[-1][0] + merged[-1][1] == b:
                merged[-1] = (merged[-1][0], merged[-1][1] + s)
            else:
                merged.append((b, s))
        self._free = merged
        return size

    @property
    def used(self) -> int:
        return sum(self._live.values())

    @property
    def live_runs(self) -> Dict[int, int]:
        return dict(self._live)
