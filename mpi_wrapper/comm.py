"""``Communicator`` -- the byte-counting collective wrapper of the reference, rebuilt on the
B200 data plane.

Same public surface and the same ``total_bytes_transferred`` accounting as the reference class
(mpi_wrapper/comm.py:4-199), but every method takes NumPy arrays *or* torch tensors (CPU or
CUDA), and the "my*" methods -- hand-rolled in the reference from blocking Send/Recv
(reduce-to-root + broadcast, :63-107) and Isend/Irecv (:110-159) -- are single device kernels
here when the ranks own GPUs:

* ``myAllreduce``  -> one-shot / two-shot / NVLS-multicast all-reduce (``algo=`` selects), the
  deterministic rank-order reduction of the reference's root loop without its 2(n-1) serial
  message latencies (README.md:141-142);
* ``myAlltoall``   -> one kernel with every peer push in flight, landing directly in ``dest``
  (no per-peer receive buffers and no copy-out, reference :136-155);
* ``myAlltoall2``  -> the same kernel walking peers pairwise in ``(rank+i) % n`` order
  (reference :186-194).

Without GPUs the same methods run over the shared-memory transport: ``Allreduce`` & co. use its
staged collectives, the "my*" variants are built from its point-to-point rings so that they
remain an independent implementation to compare against (mpi-test.py:40-98,178-239).
"""
from __future__ import annotations

import numpy as np
from mpi4py import MPI

import time as _time

from collective_communication_mpi_b200 import _native as _N
from collective_communication_mpi_b200.mpi import core as _core
from collective_communication_mpi_b200.utils.timing import OpStats as _OpStats

_ALGOS = {"auto": _N.AR_AUTO, "oneshot": _N.AR_ONESHOT, "twoshot": _N.AR_TWOSHOT, "nvls": _N.AR_NVLS,
          "multicast": _N.AR_NVLS}


def _nbytes(a) -> int:
    if hasattr(a, "element_size"):
        return a.element_size() * a.numel()
    return a.itemsize * a.size


def _count(a) -> int:
    return a.numel() if hasattr(a, "numel") else a.size


class Communicator(object):
    def __init__(self, comm: MPI.Comm):
        self.comm = comm
        self.total_bytes_transferred = 0
        # beyond the reference's single byte counter: per-op calls / payload bytes / wall seconds
        self.stats = _OpStats()

    # ---- passthroughs (reference :9-16) -------------------------------------------------
    def Get_size(self):
        return self.comm.Get_size()

    def Get_rank(self):
        return self.comm.Get_rank()

    def Barrier(self):
        return self.comm.Barrier()

    # ---- library collectives with accounting (reference :18-61) -------------------------
    def Allreduce(self, src_array, dest_array, op=MPI.SUM):
        assert _count(src_array) == _count(dest_array)
        self._account(2 * _nbytes(src_array))
        t0 = _time.perf_counter()
        self.comm.Allreduce(src_array, dest_array, op)
        self.stats.add("Allreduce", _nbytes(src_array), _time.perf_counter() - t0)

    def Allgather(self, src_array, dest_array):
        self._account(_nbytes(src_array) + _nbytes(dest_array))
        t0 = _time.perf_counter()
        self.comm.Allgather(src_array, dest_array)
        self.stats.add("Allgather", _nbytes(dest_array), _time.perf_counter() - t0)

    def Reduce_scatter(self, src_array, dest_array, op=MPI.SUM):
        self._account(_nbytes(src_array) + _nbytes(dest_array))
        t0 = _time.perf_counter()
        self.comm.Reduce_scatter_block(src_array, dest_array, op)
        self.stats.add("Reduce_scatter", _nbytes(src_array), _time.perf_counter() - t0)

    def Split(self, key, color):
        # (key, color) order is the reference's (comm.py:38), the reverse of MPI's
        return __class__(self.comm.Split(key=key, color=color))

    def Alltoall(self, src_array, dest_array):
        n = self.comm.Get_size()
        assert _count(src_array) % n == 0, "src_array size must be divisible by the number of processes"
        assert _count(dest_array) % n == 0, "dest_array size must be divisible by the number of processes"
        self._account(_nbytes(src_array) // n + _nbytes(dest_array) // n)
        t0 = _time.perf_counter()
        self.comm.Alltoall(src_array, dest_array)
        self.stats.add("Alltoall", _nbytes(src_array), _time.perf_counter() - t0)

    # ---- the hand-rolled collectives, re-designed (reference :63-199) -------------------
    def myAllreduce(self, src_array, dest_array, op=MPI.SUM, algo: str = "auto"):
        """All-reduce that does not go through the library ``Allreduce`` path.

        Device plane: one kernel; ``algo`` in {"auto", "oneshot", "twoshot", "nvls"}.
        Host plane: gather-to-root over the p2p rings, rank-ordered fold, broadcast.
        Accounting as the reference: root ``2*B*(n-1)``, others ``2*B`` (:101,107)."""
        if op not in (MPI.SUM, MPI.MIN, MPI.MAX):
            raise NotImplementedError("Only MPI.SUM, MPI.MIN, and MPI.MAX are supported.")
        rank, n = self.comm.Get_rank(), self.comm.Get_size()
        nbytes = _nbytes(src_array)
        if isinstance(self.comm, _core.Comm):
            s, r = _core._Buf(src_array, False), _core._Buf(dest_array, True)
            if n == 1:
                _core._copy(r, s)
            elif self.comm._on_device(s, r):
                _my_allreduce_device(self.comm, s, r, op.code, _ALGOS[algo])
            else:
                _my_allreduce_p2p(self.comm, src_array, dest_array, op)
        else:
            _my_allreduce_p2p(self.comm, src_array, dest_array, op)
        self.total_bytes_transferred += 2 * nbytes * (n - 1) if rank == 0 else 2 * nbytes

    def myAlltoall(self, src_array, dest_array):
        """All-to-all with every transfer in flight at once (reference :110-159)."""
        self._my_alltoall(src_array, dest_array, staggered=False)

    def myAlltoall2(self, src_array, dest_array):
        """All-to-all walking the peers pairwise (reference :162-199)."""
        self._my_alltoall(src_array, dest_array, staggered=True)

    def _my_alltoall(self, src_array, dest_array, staggered: bool):
        n = self.comm.Get_size()
        seg = _count(src_array) // n
        seg_bytes = _nbytes(src_array) // _count(src_array) * seg if _count(src_array) else 0
        if isinstance(self.comm, _core.Comm):
            s, r = _core._Buf(src_array, False), _core._Buf(dest_array, True)
            if n == 1:
                _core._copy(r, s)
            elif self.comm._on_device(s, r):
                _core._dev_alltoall(self.comm._dev, s, r, n, staggered=staggered)
            else:
                _my_alltoall_p2p(self.comm, src_array, dest_array, seg, staggered)
        else:
            _my_alltoall_p2p(self.comm, src_array, dest_array, seg, staggered)
        self.total_bytes_transferred += 2 * seg_bytes * (n - 1)

    # ---- what func_impl calls on its communicator (reference func_impl.py:89,107,184) ----
    def allgather(self, obj):
        return self.comm.allgather(obj)

    def alltoall(self, objs):
        return self.comm.alltoall(objs)

    def gather_last_axis(self, x):
        out = self.comm.gather_last_axis(x)
        self._account(_nbytes(x) + _nbytes(out))
        return out

    def reduce_scatter_last_axis(self, x, op=MPI.SUM):
        out = self.comm.reduce_scatter_last_axis(x, op)
        self._account(_nbytes(x) + _nbytes(out))
        return out

    def _account(self, per_peer_bytes: int):
        self.total_bytes_transferred += per_peer_bytes * (self.comm.Get_size() - 1)


# -------------------------------------------------------------------------------------------------
# device plane
# -------------------------------------------------------------------------------------------------
def _my_allreduce_device(comm, s, r, op_code: int, algo: int):
    from collective_communication_mpi_b200.ops import collectives as CO
    dev = comm._dev
    sin, sout, stmp, cap = dev.STAGE
    if algo == _N.AR_NVLS and not dev.has_mc:
        raise RuntimeError("algo='nvls' needs NVLS multicast, which this communicator does not have")
    dev.agree(s, r)     # host rendezvous + "same heap offset on every member?" (zero-copy only then)
    es = s.itemsize
    step = cap // es * es
    for lo in range(0, s.nbytes, step):
        nb = min(step, s.nbytes - lo)
        src = _core._stage_in(dev, s, sin, lo, nb)
        dst = _core._out_offset(dev, r, sout, lo)
        CO.allreduce(dev.comms, s.cc, op_code, src, dst, nb // es, tmp_off=stmp, algo=algo)
        _core._stage_out(dev, r, dst, lo, nb)
    _core._finish(r)


# -------------------------------------------------------------------------------------------------
# host plane: the "my*" collectives from point-to-point messages
# -------------------------------------------------------------------------------------------------
def _flat_np(a):
    if hasattr(a, "numpy") and not isinstance(a, np.ndarray):
        a = a.numpy()
    return a.reshape(-1)


def _my_allreduce_p2p(comm, src_array, dest_array, op):
    rank, n = comm.Get_rank(), comm.Get_size()
    src, dst = _flat_np(src_array), _flat_np(dest_array)
    if rank != 0:
        comm.Send(src, dest=0)
        comm.Recv(dst, source=0)
        return
    # all receives are posted up front; the fold still runs in rank order 1..n-1, which keeps
    # the result identical to the reference's sequential loop
    inbox = np.empty((n - 1,) + src.shape, dtype=src.dtype)
    reqs = [comm.Irecv(inbox[i - 1], source=i) for i in range(1, n)]
    np.copyto(dst, src)
    fold = {MPI.SUM: np.add, MPI.MIN: np.minimum, MPI.MAX: np.maximum}[op]
    for i in range(1, n):
        reqs[i - 1].Wait()
        fold(dst, inbox[i - 1], out=dst)
    MPI.Request.Waitall([comm.Isend(dst, dest=i) for i in range(1, n)])


def _my_alltoall_p2p(comm, src_array, dest_array, seg: int, staggered: bool):
    rank, n = comm.Get_rank(), comm.Get_size()
    src, dst = _flat_np(src_array), _flat_np(dest_array)
    cut = lambda a, i: a[i * seg:(i + 1) * seg]  # noqa: E731
    np.copyto(cut(dst, rank), cut(src, rank))
    if staggered:
        for i in range(1, n):
            to, frm = (rank + i) % n, (rank - i) % n
            comm.Sendrecv(cut(src, to), dest=to, sendtag=rank, recvbuf=cut(dst, frm), source=frm, recvtag=frm)
        return
    reqs = [comm.Irecv(cut(dst, i), source=i) for i in range(n) if i != rank]
    reqs += [comm.Isend(cut(src, i), dest=i) for i in range(n) if i != rank]
    MPI.Request.Waitall(reqs)
