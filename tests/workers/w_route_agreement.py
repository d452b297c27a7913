"""The zero-copy route of the front-end device collectives is taken only when EVERY member passed a
tensor at the same offset of the symmetric heap (``_DeviceSide.agree``, mpi/core.py; ADVICE round 1:
a kernel applies one offset to every peer's heap). The decision is host-side logic over the host
plane's small all-gather, so it is checked here on CPU ranks: the heap is replaced by a stand-in that
reports whatever offset a scenario prescribes for this rank."""
import ctypes as C
from types import SimpleNamespace

import numpy as np

from mpi4py import MPI
from collective_communication_mpi_b200.mpi.core import _DeviceSide

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()


class FakeHeap:
    emulated = False

    @staticmethod
    def offset_of(t):
        return t.off


def side(comm):
    d = object.__new__(_DeviceSide)
    d.world, d.heap = comm._world, FakeHeap()
    d._members, d._me, d._cid = list(comm._members), comm._me, comm._cid
    d._members_c = (C.c_int * len(comm._members))(*comm._members)
    d._route_send = np.zeros(2, dtype=np.int64)
    d._route_recv = np.zeros(2 * len(comm._members), dtype=np.int64)
    d.alias = (False, False)
    return d


def buf(off, cuda=True):
    return SimpleNamespace(cuda=cuda, flat=SimpleNamespace(off=off))


def case(d, s, r, want, why):
    d.agree(s, r)
    assert d.alias == want, f"rank {rank}: {why}: alias = {d.alias}, expected {want}"


d = side(world)
case(d, buf(4096), buf(8192), (True, True), "same offsets everywhere")
case(d, buf(4096 + 256 * rank), buf(8192), (False, True), "rank-dependent slice of a symmetric tensor as the source")
case(d, buf(4096), buf(8192 + 512 * rank), (True, False), "rank-dependent destination")
case(d, buf(None if rank == n - 1 else 4096), buf(8192), (False, True), "one member passed an ordinary CUDA tensor")
case(d, buf(4096, cuda=(rank != 0)), buf(8192), (False, True), "one member passed a host array")
case(d, buf(4104), buf(8192), (False, True), "same offset but not 16-byte aligned")
case(d, buf(0), buf(16), (True, True), "offset 0 is a valid offset")
# a sub-communicator keeps its own agreement (other id, other members), concurrently with its sibling
half = world.Split(color=rank % 2, key=rank)
dh = side(half)
many = half.Get_size() > 1       # a group of one has nobody to agree with and never takes the route
case(dh, buf(1024 * (rank % 2 + 1)), buf(64), (many, many), "siblings agree within themselves on different offsets")
case(dh, buf(1024 + 16 * half.Get_rank()), buf(64), (False, many), "rank-dependent inside a sub-communicator")
world.Barrier()
if rank == 0:
    print(f"ROUTE AGREEMENT OK ranks={n}")
