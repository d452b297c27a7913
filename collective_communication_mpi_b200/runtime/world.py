"""Process bootstrap: who am I, who else is there, and how do we reach each other.

Replaces what ``mpirun`` + ``MPI_Init`` give the reference (README.md:52, mpi-test.py:16): ranks
are ordinary OS processes started by ``python -m collective_communication_mpi_b200.launch`` or
by ``torchrun`` (both export RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT).

Two planes:

* control plane + host data plane: the shared-memory transport (csrc/host_shm.cpp). Always
  attached when there is more than one rank; small pickled messages (Split colours, shapes)
  and all NumPy-array traffic on a machine without GPUs go through it.
* device data plane: the symmetric heap (csrc/runtime.cpp), created lazily the first time a
  device transport is needed, one GPU per rank (device = LOCAL_RANK).

Single node only -- one NVSwitch domain / one shared-memory segment -- like the reference,
whose "distributed" runs are ``mpirun -n 8`` on one host (README.md:9,41).
"""
from __future__ import annotations

import atexit
import ctypes as C
import os
import pickle
import time
import uuid
from typing import List, Optional

import numpy as np

from .. import _native as N


class World:
    """Singleton describing this process's place in the job."""

    _instance: Optional["World"] = None

    def __init__(self):
        env = os.environ
        self.rank = int(env.get("RANK", env.get("CCMPI_RANK", "0")))
        self.size = int(env.get("WORLD_SIZE", env.get("CCMPI_WORLD_SIZE", "1")))
        self.local_rank = int(env.get("LOCAL_RANK", str(self.rank)))
        job = env.get("CCMPI_JOB")
        if not job:
            # torchrun children share a parent; that and the rendezvous port name the job
            job = f"{env.get('MASTER_PORT', '0')}-{os.getppid()}" if self.size > 1 else uuid.uuid4().hex[:12]
        self.job = job
        self.timeout_s = float(env.get("CCMPI_TIMEOUT", "120"))
        self.host = None           # host transport handle (int) or None
        self._hostlib = None
        self.heap = None           # SymmetricHeap or None
        self._device_tried = False
        self._device_ok = {}
        self.numa_node: Optional[int] = None
        self.device_error: Optional[str] = None
        self.mc_supported = False
        self.stage = None
        self.sm_count = 148
        self._next_comm_id = 1     # 0 is the world
        self._closed = False
        if self.size > 1:
            self._attach_host()
        atexit.register(self.close)

    # ---- singleton ----------------------------------------------------------------------
    @classmethod
    def get(cls) -> "World":
        if cls._instance is None:
            cls._instance = World()
        return cls._instance

    # ---- host plane ---------------------------------------------------------------------
    def _attach_host(self):
        lib = N.host_lib()
        stage = int(os.environ.get("CCMPI_HOST_STAGE_MB", "8")) << 20
        ring = int(os.environ.get("CCMPI_HOST_RING_KB", "1024")) << 10
        h = lib.ccmpi_host_attach(self.job.encode(), self.rank, self.size, stage, ring, self.timeout_s)
        N.check_host(h, "ccmpi_host_attach")
        self._hostlib = lib
        self.host = h
        self._world_members = (C.c_int * self.size)(*range(self.size))
        # everyone is attached before anyone proceeds (and before rank 0 could ever unlink)
        N.check_host(lib.ccmpi_host_barrier(h, 0, self._world_members, self.size, self.rank), "initial barrier")

    def host_barrier(self, slot: int, members, me: int):
        if len(members) > 1:
            arr = (C.c_int * len(members))(*members)
            N.check_host(self._hostlib.ccmpi_host_barrier(self.host, slot, arr, len(members), me), "barrier")

    def host_allgather_bytes(self, slot: int, members, me: int, payload: bytes) -> List[bytes]:
        """All-gather of variable-length byte strings over the host plane (control messages,
        and the pickle path of ``comm.allgather`` -- reference func_impl.py:89)."""
        n = len(members)
        if n == 1:
            return [payload]
        arr = (C.c_int * n)(*members)
        lib, h = self._hostlib, self.host
        lens = np.zeros(n, dtype=np.int64)
        mine = np.array([len(payload)], dtype=np.int64)
        N.check_host(lib.ccmpi_host_allgather(h, slot, arr, n, me, mine.ctypes.data, lens.ctypes.data, 8), "allgather(len)")
        width = int(lens.max())
        if width == 0:
            return [b""] * n
        send = np.zeros(width, dtype=np.uint8)
        send[:len(payload)] = np.frombuffer(payload, dtype=np.uint8)
        recv = np.empty(n * width, dtype=np.uint8)
        N.check_host(lib.ccmpi_host_allgather(h, slot, arr, n, me, send.ctypes.data, recv.ctypes.data, width), "allgather(data)")
        return [recv[i * width:i * width + int(lens[i])].tobytes() for i in range(n)]

    def host_allgather_obj(self, slot: int, members, me: int, obj):
        return [pickle.loads(b) for b in self.host_allgather_bytes(slot, members, me, pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL))]

    # ---- device plane -------------------------------------------------------------------
    def device_available(self) -> bool:
        """True when every rank of this node can have its own GPU."""
        mode = os.environ.get("CCMPI_TRANSPORT", "auto").lower()
        if mode == "host":
            return False
        cached = self._device_ok.get(mode)      # asked on every collective call: probe the driver once
        if cached is not None:
            return cached
        try:
            import torch
            ok = torch.cuda.is_available() and torch.cuda.device_count() > self.local_rank and \
                torch.cuda.device_count() >= min(self.size, N.MAX_RANKS) and self.size <= N.MAX_RANKS
        except Exception:  # noqa: BLE001
            ok = False
        if mode == "device" and not ok:
            raise RuntimeError("CCMPI_TRANSPORT=device but there is no GPU per rank on this node")
        self._device_ok[mode] = ok
        return ok

    def ensure_heap(self):
        """Create + exchange the symmetric heap (collective over the world). Idempotent."""
        if self.heap is not None or self._device_tried:
            return self.heap
        self._device_tried = True
        from .symm import SymmetricHeap
        import torch
        lib = N.device_lib()
        dev = self.local_rank
        torch.cuda.set_device(dev)
        if os.environ.get("CCMPI_NUMA_BIND") == "1":      # opt-in: run next to the GPU (utils/numa.py)
            from ..utils.numa import bind_to_gpu_node
            self.numa_node = bind_to_gpu_node(dev)
        torch.zeros(1, device=f"cuda:{dev}")  # primary context
        probe = (C.c_int * 8)()
        N.check(lib.ccmpi_probe(dev, probe), "ccmpi_probe")
        self.sm_count = int(probe[3]) or 148
        mb = int(os.environ.get("CCMPI_HEAP_MB", "2048"))
        nbytes = mb << 20
        if self.size == 1:
            # one rank: a plain device buffer behaves like a 1-member heap
            h = lib.ccmpi_heap_create_emulated(dev, nbytes, 1)
            N.check(h, "ccmpi_heap_create_emulated")
            self.heap = SymmetricHeap(h, 1, 0, dev, emulated=True)
            self._alloc_stage()
            return self.heap
        h = lib.ccmpi_heap_create(dev, nbytes)
        N.check(h, "ccmpi_heap_create")
        job = self.job.encode()
        N.check(lib.ccmpi_heap_listen(h, job, self.rank), "ccmpi_heap_listen")
        self.host_barrier(0, list(range(self.size)), self.rank)  # every rank listens
        N.check(lib.ccmpi_heap_connect_all(h, job, self.rank, self.size), "ccmpi_heap_connect_all")
        self.host_barrier(0, list(range(self.size)), self.rank)  # every mapping exists
        self.heap = SymmetricHeap(h, self.size, self.rank, dev)
        self.mc_supported = bool(probe[4]) and os.environ.get("CCMPI_NVLS", "1") != "0"
        self._alloc_stage()
        return self.heap

    def _alloc_stage(self):
        """Staging arena for host arrays / non-symmetric tensors: in, out, scratch. Allocated
        here, world-collectively, so the offsets agree on every rank."""
        nb = int(os.environ.get("CCMPI_STAGE_MB", "64")) << 20
        self.stage = (self.heap.alloc(nb), self.heap.alloc(nb), self.heap.alloc(nb), nb)

    def make_multicast(self, members: List[int], me: int, slot: int, tag: int) -> int:
        """Multicast alias of the members' heaps; returns the device pointer or 0. Collective
        over ``members`` (world ranks)."""
        if not self.mc_supported or len(members) < 2 or self.heap is None or self.heap.emulated:
            return 0
        lib = N.device_lib()
        arr = (C.c_int * len(members))(*members)
        g = lib.ccmpi_mc_create(self.heap.handle, self.job.encode(), self.rank, arr, len(members), tag)
        ok = g >= 0
        # agree on the outcome: a member that failed must not leave the others bound to a
        # multicast object it is not part of
        flags = self.host_allgather_obj(slot, members, me, bool(ok))
        if not all(flags):
            if not ok:
                self.device_error = N.device_lib().ccmpi_last_error().decode(errors="replace")
            self.mc_supported = False
            return 0
        rc = lib.ccmpi_mc_bind(self.heap.handle, g)
        flags = self.host_allgather_obj(slot, members, me, rc == 0)
        if not all(flags):
            self.mc_supported = False
            return 0
        return int(lib.ccmpi_mc_ptr(self.heap.handle, g) or 0)

    # ---- communicator ids ---------------------------------------------------------------
    def reserve_comm_ids(self, count: int) -> int:
        """``count`` fresh communicator ids from the job-wide counter in the shared segment (a
        process-local counter when the job is a single process). Called by ONE member of the
        parent communicator, which then tells the others: ids derived from the members' own
        state would collide between groups that have no rank in common."""
        if self.host is None:
            base = self._next_comm_id
            self._next_comm_id += count
            return base
        base = self._hostlib.ccmpi_host_reserve_comm_ids(self.host, count)
        N.check_host(min(base, 0), "reserve_comm_ids")
        return base

    # ---- teardown -----------------------------------------------------------------------
    def close(self):
        if self._closed:
            return
        self._closed = True
        try:
            if self.heap is not None:
                try:
                    import torch
                    torch.cuda.synchronize()
                except Exception:  # noqa: BLE001
                    pass
                if self.host is not None:
                    # nobody unmaps while a peer may still touch the segment
                    try:
                        self.host_barrier(0, list(range(self.size)), self.rank)
                    except Exception:  # noqa: BLE001
                        pass
                self.heap.close()
                self.heap = None
        finally:
            if self.host is not None:
                self._hostlib.ccmpi_host_detach(self.host, 1 if self.rank == 0 else 0)
                self.host = None


def wtime() -> float:
    """MPI.Wtime (mpi-test.py:60)."""
    return time.perf_counter()
