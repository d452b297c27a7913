"""Point-to-point transfers of CUDA tensors over NVLink (the default on the device plane; ``CCMPI_P2P=host`` stages
them through the host rings instead).

The reference moves everything with ``Send`` / ``Recv`` / ``Isend`` / ``Irecv`` / ``Sendrecv`` on
host arrays (mpi_wrapper/comm.py:86,99,104-105,140,147,192). The default path here keeps that
behaviour for CUDA tensors by staging them through the host rings. This module is the direct
route ("put-with-signal mailbox ... CUDA / memcpy"):

* **send**: the payload is copied into a block of the sender's symmetric heap (device to device),
  the stream is synchronised, and a 32-byte *descriptor* -- magic, heap offset, byte count,
  sequence number -- goes through the host ring with the caller's ``(communicator, tag)``; ``Send``
  returns as soon as the descriptor is queued (buffered-send semantics, so the "everyone sends,
  then everyone receives" exchange cannot deadlock);
* **receive**: when the descriptor arrives the receiver copies the payload straight out of the
  sender's heap through its peer mapping -- the bytes cross NVLink once, no host bounce -- waits
  for the copy, and bumps the pair's ack counter in the shared-memory segment
  (``ccmpi_host_ack_post``);
* **reclaim**: the sender frees staging blocks whose sequence number the counter has passed,
  lazily, on its next point-to-point call (or when the heap runs short).

Ordering, matching and error reporting are those of the host rings (strict FIFO per pair, tag and
communicator checked). A receiver may post a NumPy / CPU buffer against a CUDA sender: it recognises
the descriptor and copies the payload out of the sender's staging block (``complete_receive_host``).

The control path (descriptors, ack counters, reclamation) runs in the CPU suite against a stand-in
heap (tests/test_p2p_device_control.py); the NVLink copy on 2 / 4 / 8 B200s (tests/test_gpu_p2p_device.py).
"""
from __future__ import annotations

import os
import time
from collections import deque
from typing import Deque, Dict, Tuple

import numpy as np

DESC_MAGIC = 0x4343_4D50_4944_4556          # "CCMPIDEV"
DESC_WORDS = 4                              # magic, offset, nbytes, sequence number
DESC_BYTES = DESC_WORDS * 8


def enabled() -> bool:
    return os.environ.get("CCMPI_P2P", "device") == "device"


def make_descriptor(off: int, nbytes: int, seq: int) -> np.ndarray:
    return np.array([DESC_MAGIC, off, nbytes, seq], dtype=np.int64)


def parse_descriptor(raw: np.ndarray, expect_bytes: int):
    """``(offset, nbytes, seq)`` if ``raw`` starts with a descriptor for ``expect_bytes`` or fewer
    bytes, else ``None`` (the sender used the host path and ``raw`` is the payload itself)."""
    if raw.nbytes < DESC_BYTES:
        return None
    w = raw[:DESC_BYTES].view(np.int64)
    if int(w[0]) != DESC_MAGIC or not 0 <= int(w[2]) <= expect_bytes:
        return None
    return int(w[1]), int(w[2]), int(w[3])


class StagingLedger:
    """Sender-side bookkeeping of staging blocks per destination: blocks are handed out in send
    order and released when the destination's ack counter has passed them."""

    def __init__(self, alloc, free, ack_count, timeout_s: float = 120.0):
        self._alloc, self._free, self._ack_count = alloc, free, ack_count
        self._sent: Dict[int, int] = {}
        self._live: Dict[int, Deque[Tuple[int, int]]] = {}
        self.timeout_s = timeout_s

    def reclaim(self, dst: int = None) -> int:
        freed = 0
        for d in ([dst] if dst is not None else list(self._live)):
            q = self._live.get(d)
            if not q:
                continue
            done = self._ack_count(d)
            while q and q[0][0] < done:
                self._free(q.popleft()[1])
                freed += 1
        return freed

    def stage(self, dst: int, nbytes: int) -> Tuple[int, int]:
        """Reserve a block for the next message to ``dst``: ``(offset, sequence number)``. Waits
        for acknowledgements when the heap is full."""
        self.reclaim(dst)
        t0 = time.monotonic()
        while True:
            try:
                off = self._alloc(max(nbytes, 16))
                break
            except MemoryError:
                if self.reclaim() == 0:
                    if time.monotonic() - t0 > self.timeout_s:
                        raise
                    time.sleep(0.0005)
        seq = self._sent.get(dst, 0)
        self._sent[dst] = seq + 1
        self._live.setdefault(dst, deque()).append((seq, off))
        return off, seq

    @property
    def outstanding(self) -> int:
        return sum(len(q) for q in self._live.values())


# ---- glue used by Comm.Isend / Comm.Irecv -------------------------------------------------------
def ledger_for(world) -> StagingLedger:
    led = getattr(world, "_p2p_ledger", None)
    if led is None:
        heap = world.ensure_heap()
        lib, h = world._hostlib, world.host

        def ack_count(dst):
            return int(lib.ccmpi_host_ack_count(h, dst))

        led = StagingLedger(heap.alloc, heap.free, ack_count, world.timeout_s)
        world._p2p_ledger = led
    return led


def stage_for_send(world, dst_world: int, flat_cuda) -> np.ndarray:
    """Copy a flat CUDA tensor into the symmetric heap and return the descriptor to send."""
    import torch
    heap = world.ensure_heap()
    nbytes = flat_cuda.numel() * flat_cuda.element_size()
    off, seq = ledger_for(world).stage(dst_world, nbytes)
    if nbytes:
        heap.view(off, (nbytes,), torch.uint8).copy_(flat_cuda.view(torch.uint8), non_blocking=True)
        torch.cuda.current_stream().synchronize()       # the peer reads this memory next
    return make_descriptor(off, nbytes, seq)


def complete_receive(world, src_world: int, raw: np.ndarray, flat_cuda) -> None:
    """``raw`` is what arrived through the ring for a CUDA receive buffer: a descriptor (pull the
    payload over NVLink, then acknowledge) or, if the sender used the host path, the payload."""
    import torch
    nbytes = flat_cuda.numel() * flat_cuda.element_size()
    d = parse_descriptor(raw, nbytes)
    if d is None:
        flat_cuda.view(torch.uint8).copy_(torch.from_numpy(np.ascontiguousarray(raw[:nbytes])))
        torch.cuda.current_stream().synchronize()
        return
    off, n, _ = d
    if n:
        heap = world.ensure_heap()
        flat_cuda.view(torch.uint8)[:n].copy_(heap.view(off, (n,), torch.uint8, rank=src_world), non_blocking=True)
        torch.cuda.current_stream().synchronize()
    rc = world._hostlib.ccmpi_host_ack_post(world.host, src_world)
    if rc < 0:
        raise RuntimeError("ccmpi_host_ack_post failed")


def complete_receive_host(world, src_world: int, raw: np.ndarray, dst_bytes: np.ndarray) -> None:
    """The same for a HOST receive buffer (``dst_bytes``: its uint8 view): MPI lets a CUDA sender meet a NumPy
    receiver, so a host receive on the device plane must also understand a descriptor."""
    nbytes = dst_bytes.nbytes
    d = parse_descriptor(raw, nbytes)
    if d is None:
        np.copyto(dst_bytes, raw[:nbytes])
        return
    import torch
    off, n, _ = d
    if n:
        heap = world.ensure_heap()
        torch.from_numpy(dst_bytes)[:n].copy_(heap.view(off, (n,), torch.uint8, rank=src_world))
        torch.cuda.current_stream().synchronize()
    rc = world._hostlib.ccmpi_host_ack_post(world.host, src_world)
    if rc < 0:
        raise RuntimeError("ccmpi_host_ack_post failed")
