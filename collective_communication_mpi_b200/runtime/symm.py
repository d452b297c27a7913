"""Symmetric heap: one equally sized device segment per rank, every segment mapped into every
process, so ``(rank, offset)`` names a buffer on any GPU and ``offset`` alone names "the same
buffer on every rank".

This is the data-plane half of what ``MPI.COMM_WORLD`` is for the reference (mpi-test.py:16,
tests/test_get_info.py:57). The native side (csrc/runtime.cpp) creates and maps segments;
this module owns the offset allocator and hands out ``torch`` tensors that alias heap memory,
so kernels can be given plain offsets and users plain tensors.

Heap layout::

    [0, PAD_REGION)            signal pads, one 64 KiB pad per communicator id
    [PAD_REGION, bytes)        allocator arena (256-byte aligned blocks)

Allocation is *collective by convention*: every rank of the world performs the same sequence
of ``alloc``/``free`` calls, which keeps offsets identical everywhere without communication.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

from .. import _native as N

MAX_COMMS = 64
PAD_REGION = MAX_COMMS * N.PAD_BYTES  # 4 MiB
ALIGN = 256


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` carrier so torch can alias a raw device pointer."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,),
            "typestr": "|u1",
            "data": (ptr, False),
            "version": 3,
            "strides": None,
        }


def _torch_dtype_size(dtype) -> int:
    import torch
    return torch.empty((), dtype=dtype).element_size()


class SymmetricHeap:
    def __init__(self, handle: int, nranks: int, rank: int, device: int, emulated: bool = False):
        self.lib = N.device_lib()
        self.handle = handle
        self.nranks = nranks
        self.rank = rank
        self.device = device
        self.emulated = emulated
        self.bytes = int(self.lib.ccmpi_heap_bytes(handle))
        self.ptrs: List[int] = []
        for r in range(nranks):
            p = self.lib.ccmpi_heap_peer_ptr(handle, r)
            if not p:
                raise N.NativeError(f"heap segment of rank {r} is not mapped")
            self.ptrs.append(int(p))
        N.apply_device_timeout(device)      # watchdog budget of the kernels' waits (CCMPI_DEVICE_TIMEOUT_S)
        self._free: List[Tuple[int, int]] = [(PAD_REGION, self.bytes - PAD_REGION)]  # (offset, size)
        self._live: Dict[int, int] = {}
        self._bytes_views: Dict[int, object] = {}
        self._closed = False

    # ---- construction -------------------------------------------------------------------
    @classmethod
    def create_emulated(cls, nranks: int, nbytes: int, device: int = 0) -> "SymmetricHeap":
        """``nranks`` plain device buffers on ONE GPU (single-GPU tests of multi-rank kernels)."""
        lib = N.device_lib()
        h = lib.ccmpi_heap_create_emulated(device, PAD_REGION + nbytes, nranks)
        N.check(h, "ccmpi_heap_create_emulated")
        return cls(h, nranks, 0, device, emulated=True)

    # ---- allocator ----------------------------------------------------------------------
    def alloc(self, nbytes: int) -> int:
        """Reserve ``nbytes`` (rounded to 256 B) at the same offset on every rank."""
        size = max(ALIGN, (int(nbytes) + ALIGN - 1) // ALIGN * ALIGN)
        for i, (off, sz) in enumerate(self._free):
            if sz >= size:
                if sz == size:
                    self._free.pop(i)
                else:
                    self._free[i] = (off + size, sz - size)
                self._live[off] = size
                return off
        raise MemoryError(
            f"symmetric heap exhausted: need {size} B, largest free block "
            f"{max((s for _, s in self._free), default=0)} B of {self.bytes} B "
            f"(raise CCMPI_HEAP_MB)")

    def free(self, off: int) -> None:
        size = self._live.pop(off)
        self._free.append((off, size))
        self._free.sort()
        merged: List[Tuple[int, int]] = []
        for o, s in self._free:
            if merged and merged[-1][0] + merged[-1][1] == o:
                merged[-1] = (merged[-1][0], merged[-1][1] + s)
            else:
                merged.append((o, s))
        self._free = merged

    # ---- tensor views -------------------------------------------------------------------
    def _bytes_view(self, rank: int):
        import torch
        v = self._bytes_views.get(rank)
        if v is None:
            v = torch.as_tensor(_RawCuda(self.ptrs[rank], self.bytes), device=f"cuda:{self.device}")
            self._bytes_views[rank] = v
        return v

    def view(self, off: int, shape, dtype, rank: Optional[int] = None):
        """A tensor aliasing ``[off, off + prod(shape)*itemsize)`` of ``rank``'s segment
        (default: this rank's own). Peer views are real peer-mapped memory: reading them moves
        data over NVLink."""
        r = self.rank if rank is None else rank
        n = int(math.prod(shape)) * _torch_dtype_size(dtype)
        if off < 0 or off + n > self.bytes:
            raise ValueError(f"view [{off}, {off + n}) outside the heap of {self.bytes} B")
        return self._bytes_view(r)[off:off + n].view(dtype).view(*shape)

    def empty(self, shape, dtype, rank: Optional[int] = None):
        """Allocate and return ``(offset, tensor)``."""
        n = int(math.prod(shape)) * _torch_dtype_size(dtype)
        off = self.alloc(n)
        return off, self.view(off, shape, dtype, rank)

    def offset_of(self, tensor) -> Optional[int]:
        """Heap offset of a tensor that lives in this rank's segment, else ``None``."""
        if not tensor.is_cuda:
            return None
        p = tensor.data_ptr()
        base = self.ptrs[self.rank]
        if base <= p < base + self.bytes:
            return p - base
        return None

    def pad_ptr(self, rank: int, comm_id: int) -> int:
        if not 0 <= comm_id < MAX_COMMS:
            raise ValueError(f"communicator id {comm_id} out of range (max {MAX_COMMS})")
        return self.ptrs[rank] + comm_id * N.PAD_BYTES

    def close(self) -> None:
        if not self._closed:
            self._closed = True
            self._bytes_views.clear()
            self.lib.ccmpi_heap_destroy(self.handle)


def make_dev_comm(heap: SymmetricHeap, members: List[int], me: int, comm_id: int, mc_ptr: int = 0) -> N.DevComm:
    """Assemble the device-side communicator for group ``members`` (heap ranks, in group order)
    as seen by group rank ``me``. Device analogue of the object ``MPI_Comm_split`` returns
    (reference comm.py:38-39)."""
    if len(members) > N.MAX_RANKS:
        raise ValueError("device communicators span at most one NVSwitch domain (8 GPUs)")
    c = N.DevComm()
    c.rank = me
    c.size = len(members)
    for i, r in enumerate(members):
        c.heap[i] = heap.ptrs[r]
        c.pad[i] = heap.pad_ptr(r, comm_id)
    c.mc_heap = mc_ptr or None
    return c
