"""Rank-side body of the collective tests. Run under the launcher; works on either data plane
(``--transport host|device``) and, with ``--cuda``, on CUDA tensors instead of NumPy arrays.
Oracles are closed-form NumPy expressions of what MPI defines for each collective."""
import os
import sys

import numpy as np

from mpi4py import MPI
from mpi_wrapper import Communicator

CUDA = "--cuda" in sys.argv
TORCH_CPU = "--torch-cpu" in sys.argv        # CPU tensors instead of NumPy arrays (same memory, other front door)
world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()

if CUDA or TORCH_CPU:
    import torch
if CUDA:
    torch.cuda.set_device(rank)


def put(a):
    if CUDA:
        return torch.from_numpy(np.ascontiguousarray(a)).cuda()
    if TORCH_CPU:
        return torch.from_numpy(np.array(a, copy=True))
    return a


def get(a):
    return a.cpu().numpy() if (CUDA or TORCH_CPU) else a


def empty(shape, dtype):
    return put(np.zeros(shape, dtype=dtype))


def check_buffer_collectives(comm, members):
    me, k = comm.Get_rank(), comm.Get_size()
    for dtype in (np.int64, np.int32, np.float32, np.float64):
        for count in (1, 7, 100, 4096, 70001):
            mine = ((np.arange(count) * 3 + members[me] * 7) % 101).astype(dtype)
            everyone = np.stack([((np.arange(count) * 3 + m * 7) % 101).astype(dtype) for m in members])
            for op, fold in ((MPI.SUM, np.sum), (MPI.MIN, np.min), (MPI.MAX, np.max)):
                out = empty(count, dtype)
                comm.Allreduce(put(mine), out, op=op)
                np.testing.assert_allclose(get(out), fold(everyone, axis=0).astype(dtype), rtol=1e-6)
    # Allgather: rank-ordered concatenation
    for per in (2, 33, 5000):
        out = empty(per * k, np.int64)
        comm.Allgather(put(np.arange(per, dtype=np.int64) + 1000 * me), out)
        np.testing.assert_array_equal(get(out), np.concatenate([np.arange(per) + 1000 * j for j in range(k)]))
    # Reduce_scatter_block: rank r keeps block r of the reduction
    for per in (2, 129):
        src = (np.arange(per * k, dtype=np.float64) + me)
        out = empty(per, np.float64)
        comm.Reduce_scatter_block(put(src), out, op=MPI.SUM)
        full = sum(np.arange(per * k, dtype=np.float64) + j for j in range(k))
        np.testing.assert_allclose(get(out), full[me * per:(me + 1) * per])
    # Alltoall: block j of rank i ends up as block i of rank j
    for seg in (1, 3, 2048):
        src = np.concatenate([np.full(seg, me * 100 + j, dtype=np.int64) for j in range(k)])
        out = empty(seg * k, np.int64)
        comm.Alltoall(put(src), out)
        np.testing.assert_array_equal(get(out), np.concatenate([np.full(seg, j * 100 + me) for j in range(k)]))


def check_wrapper_accounting():
    """Byte-count oracles of SURVEY 2.2 (reference comm.py:21,27-28,34-35,58-59,101,107,159)."""
    c = Communicator(world)
    a, out = put(np.arange(100, dtype=np.int64)), empty(100, np.int64)
    c.Allreduce(a, out, op=MPI.MIN)
    assert c.total_bytes_transferred == 800 * 2 * (n - 1)
    c = Communicator(world)
    c.Allgather(put(np.arange(2, dtype=np.int64)), empty(2 * n, np.int64))
    assert c.total_bytes_transferred == (16 + 16 * n) * (n - 1)
    c = Communicator(world)
    c.Reduce_scatter(put(np.arange(2 * n, dtype=np.int64)), empty(2, np.int64), op=MPI.MIN)
    assert c.total_bytes_transferred == (16 * n + 16) * (n - 1)
    c = Communicator(world)
    c.Alltoall(put(np.arange(n, dtype=np.int64)), empty(n, np.int64))
    assert c.total_bytes_transferred == 16 * (n - 1)
    c = Communicator(world)
    lib, mine = empty(100, np.int64), empty(100, np.int64)
    src = put(np.random.RandomState(rank).randint(0, 100, 100).astype(np.int64))
    for op in (MPI.MIN, MPI.MAX, MPI.SUM):
        c.comm.Allreduce(src, lib, op=op)
        c.myAllreduce(src, mine, op=op)
        np.testing.assert_array_equal(get(lib), get(mine))
    assert c.total_bytes_transferred == 3 * (2 * 800 * (n - 1) if rank == 0 else 1600)
    c = Communicator(world)
    s = put(np.array([rank * 100 + i for i in range(n)], dtype=np.int64))
    d0, d1, d2 = empty(n, np.int64), empty(n, np.int64), empty(n, np.int64)
    c.comm.Alltoall(s, d0)
    c.myAlltoall(s, d1)
    c.myAlltoall2(s, d2)
    np.testing.assert_array_equal(get(d0), get(d1))
    np.testing.assert_array_equal(get(d0), get(d2))
    assert c.total_bytes_transferred == 2 * (2 * 8 * (n - 1))
    try:
        c.myAllreduce(src, mine, op=MPI.PROD)
        raise AssertionError("PROD must be rejected")
    except NotImplementedError:
        pass
    sub = c.Split(key=rank, color=rank % 2)
    assert isinstance(sub, Communicator) and sub.total_bytes_transferred == 0


def check_point_to_point():
    if n < 2:
        return
    nxt, prv = (rank + 1) % n, (rank - 1) % n
    got = empty(1000, np.float64)
    world.Sendrecv(put(np.full(1000, float(rank))), dest=nxt, sendtag=rank, recvbuf=got, source=prv, recvtag=prv)
    np.testing.assert_array_equal(get(got), np.full(1000, float(prv)))
    big = np.arange(400_000, dtype=np.float64) + rank  # larger than a ring: must stream
    reqs, inbox = [], {}
    for j in range(n):
        if j != rank:
            inbox[j] = np.empty_like(big)
            reqs.append(world.Irecv(inbox[j], source=j))
    reqs += [world.Isend(big, dest=j) for j in range(n) if j != rank]
    MPI.Request.Waitall(reqs)
    for j, buf in inbox.items():
        np.testing.assert_array_equal(buf, np.arange(400_000) + j)


def check_device_put_signal():
    """put-with-signal mailbox kernels (the device-side Send/Recv primitive): ring exchange."""
    from collective_communication_mpi_b200.ops import collectives as CO
    dev = world.device_comm()
    heap = dev.heap
    src = world.symm_empty(4096, torch.float32)
    dst = world.symm_empty(4096, torch.float32)
    nxt, prv = (rank + 1) % n, (rank - 1) % n
    base = 12000                                   # flag words nobody else uses in this test
    for it in range(3):
        src.fill_(float(rank * 10 + it))
        torch.cuda.synchronize()
        world.Barrier()
        CO.put_signal(dev.comm, nxt, heap.offset_of(src), heap.offset_of(dst), src.numel() * 4, base + 2 * rank)
        CO.wait_signal(dev.comm, prv, base + 2 * prv)
        torch.cuda.synchronize()
        assert bool((dst == float(prv * 10 + it)).all()), (rank, it, dst[:4].tolist())
        world.Barrier()


check_buffer_collectives(world, list(range(n)))
if CUDA and n >= 2:
    check_device_put_signal()
if n >= 4:
    # sub-communicators: interleaved groups (the dp_comm pattern: stride = number of groups)
    sub = world.Split(color=rank % 2, key=rank)
    check_buffer_collectives(sub, [r for r in range(n) if r % 2 == rank % 2])
    assert world.Split(color=rank % 2, key=rank) is sub, "identical splits are cached, not re-created"
    rev = world.Split(color=0, key=-rank)
    assert rev.Get_rank() == n - 1 - rank, "key orders the new ranks"
check_wrapper_accounting()
check_point_to_point()
assert world.bcast({"k": rank} if rank == 0 else None) == {"k": 0}
assert world.allreduce(rank, op=MPI.SUM) == n * (n - 1) // 2
world.Barrier()
if rank == 0:
    print(f"COLLECTIVES OK ranks={n} cuda={CUDA}" + (" torch-cpu" if TORCH_CPU else ""))
