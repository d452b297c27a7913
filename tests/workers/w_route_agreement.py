"""The zero-copy route of the front-end device collectives is taken only when EVERY member passed a
tensor at the same offset of the symmetric heap (``_DeviceSide.agree``, mpi/core.py; ADVICE round 1:
a kernel applies one offset to every peer's heap). The decision is host-side logic over the host
plane's small all-gather, so it is checked here on CPU ranks: the heap is replaced by a stand-in that
reports whatever offset a scenario prescribes for this rank."""
import ctypes as C
import os
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
    d._route_send = np.zeros(3, dtype=np.int64)
    d._route_recv = np.zeros(3 * len(comm._members), dtype=np.int64)
    d.alias = (False, False)
    return d


def buf(off, cuda=True, nbytes=800):
    return SimpleNamespace(cuda=cuda, nbytes=nbytes, flat=SimpleNamespace(off=off))


def case(d, s, r, want, why):
    host = d.agree(s, r)
    assert not host, f"rank {rank}: {why}: host route taken without allow_host"
    assert d.alias == want, f"rank {rank}: {why}: alias = {d.alias}, expected {want}"


def host_case(d, s, r, want, why, allow=True):
    got = d.agree(s, r, allow_host=allow)
    assert got == want, f"rank {rank}: {why}: host route = {got}, expected {want}"


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
# (3) small host payloads stay on the host plane -- only when EVERY member holds one
from collective_communication_mpi_b200.mpi.core import host_route_max
assert host_route_max() == 64 << 10
h = lambda nbytes=800: buf(None, cuda=False, nbytes=nbytes)      # noqa: E731
host_case(d, h(), h(), True, "800-byte NumPy arrays everywhere (the reference's mpi-test.py payload)")
host_case(d, h(), h(), False, "library call that did not ask for the route", allow=False)
host_case(d, h(64 << 10), h(64 << 10), True, "exactly the limit")
host_case(d, h(800), h((64 << 10) + 8), False, "receive buffer above the limit (an all-gather's result)")
host_case(d, buf(4096) if rank == 0 else h(), h(), False, "one member holds a CUDA tensor: everybody runs the kernel")
host_case(d, h(), buf(8192) if rank == n - 1 else h(), False, "one member receives into a CUDA tensor")
os.environ["CCMPI_HOST_ROUTE_MAX"] = "0"
host_case(d, h(), h(), False, "route switched off")
os.environ["CCMPI_HOST_ROUTE_MAX"] = "1048576"
host_case(d, h(1 << 20), h(1 << 20), True, "limit raised by the environment")
del os.environ["CCMPI_HOST_ROUTE_MAX"]
# ... and the library collectives really fall through to the host plane then. The device plane is faked
# "up" (no GPU here); any attempt to stage or launch would fail loudly, so passing means the host route ran.
from collective_communication_mpi_b200.mpi import core
world._dev = d
core.Comm._on_device = lambda self, *bufs: True
x = np.arange(100, dtype=np.int64) + rank
out = np.empty_like(x)
world.Allreduce(x, out)
assert np.array_equal(out, n * np.arange(100) + n * (n - 1) // 2)
world.Allreduce(x, out, op=MPI.MAX)
assert np.array_equal(out, np.arange(100) + n - 1)
g = np.empty(100 * n, dtype=np.int64)
world.Allgather(x, g)
assert np.array_equal(g.reshape(n, 100), np.arange(100)[None, :] + np.arange(n)[:, None])
rs = np.empty(25, dtype=np.int64)
world.Reduce_scatter_block(np.tile(x, n)[:25 * n] * 0 + np.arange(25 * n) + rank, rs)
assert np.array_equal(rs, n * (np.arange(25) + 25 * rank) + n * (n - 1) // 2)
a2a = np.empty(4 * n, dtype=np.int64)
world.Alltoall(np.repeat(np.arange(n), 4) * 100 + rank, a2a)
assert np.array_equal(a2a, np.repeat(rank * 100 + np.arange(n), 4))
sh = np.full((2, 3, 5), float(rank))                     # the activation layout of func_impl.py
full = world.gather_last_axis(sh)
assert full.shape == (2, 3, 5 * n) and np.array_equal(full[0, 0], np.repeat(np.arange(n, dtype=float), 5))
part = world.reduce_scatter_last_axis(np.ones((2, 3, 4 * n)) * (rank + 1))
assert part.shape == (2, 3, 4) and np.all(part == n * (n + 1) / 2)
world.Barrier()
if rank == 0:
    print(f"ROUTE AGREEMENT OK ranks={n}")
