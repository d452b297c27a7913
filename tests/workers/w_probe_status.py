"""``MPI.Status``, ``Probe`` / ``Iprobe`` and receives from ``MPI.ANY_SOURCE`` on the host plane: a
master/worker exchange (every worker reports to rank 0 in its own time, rank 0 takes the reports as they
come and answers the sender), message sizes read from the status, tags matched, FIFO per pair kept."""
import time

import numpy as np

from mpi4py import MPI

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
assert n >= 3

# 1. master / worker with ANY_SOURCE + Status (buffer API)
if rank == 0:
    seen = {}
    for _ in range(n - 1):
        st = MPI.Status()
        buf = np.full(16, -1, dtype=np.int64)
        world.Recv(buf, source=MPI.ANY_SOURCE, tag=MPI.ANY_TAG, status=st)
        src = st.Get_source()
        assert st.Get_tag() == 100 + src, (src, st.Get_tag())
        cnt = st.Get_count(MPI.INT64_T)
        assert cnt == src + 1 and st.Get_count() == 8 * cnt, (src, cnt)       # shorter than the buffer: MPI allows it
        assert np.array_equal(buf[:cnt], np.arange(cnt) + 1000 * src) and buf[cnt] == -1
        seen[src] = cnt
        world.Send(np.array([cnt * 2], dtype=np.int64), dest=src, tag=7)
    assert sorted(seen) == list(range(1, n))
else:
    time.sleep(0.01 * ((rank * 7) % n))                                          # arrive in scrambled order
    world.Send(np.arange(rank + 1, dtype=np.int64) + 1000 * rank, dest=0, tag=100 + rank)
    ans = np.empty(1, dtype=np.int64)
    world.Recv(ans, source=0, tag=7)
    assert ans[0] == 2 * (rank + 1)
world.Barrier()

# 2. Probe tells the size before the buffer exists; Iprobe is false when nothing is there
if rank == 1:
    world.Send(np.arange(37, dtype=np.float64), dest=2, tag=5)
elif rank == 2:
    st = MPI.Status()
    assert world.Probe(source=1, tag=5, status=st)
    assert (st.Get_source(), st.Get_tag(), st.Get_count(MPI.DOUBLE)) == (1, 5, 37)
    got = np.empty(st.Get_count(MPI.DOUBLE))
    world.Recv(got, source=1, tag=5)
    assert np.array_equal(got, np.arange(37.0))
    assert not world.Iprobe(source=1)
    assert not world.Iprobe()                                                     # ANY_SOURCE: nobody has anything for us
world.Barrier()

# 3. tags: a probe for another tag does not match the message at the head of the ring (FIFO per pair, as a
#    posted receive with that tag would not either); with the right tag it does
if rank == 0:
    world.Send(np.array([1], dtype=np.int32), dest=1, tag=11)
    world.Send(np.array([2], dtype=np.int32), dest=1, tag=22)
elif rank == 1:
    st = MPI.Status()
    world.Probe(source=0, tag=MPI.ANY_TAG, status=st)
    assert st.Get_tag() == 11
    assert not world.Iprobe(source=0, tag=22)
    a, b = np.empty(1, dtype=np.int32), np.empty(1, dtype=np.int32)
    world.Recv(a, source=0, tag=11)
    assert world.Iprobe(source=0, tag=22, status=st) and st.Get_count(MPI.INT) == 1
    world.Recv(b, source=0, tag=22)
    assert (a[0], b[0]) == (1, 2)
world.Barrier()

# 4. a sub-communicator's ANY_SOURCE only sees its own members and ids, ranks are group ranks
sub = world.Split(color=rank % 2, key=rank)
if sub.Get_size() >= 2:
    if sub.Get_rank() == 0:
        st = MPI.Status()
        for _ in range(sub.Get_size() - 1):
            obj = sub.recv(source=MPI.ANY_SOURCE, status=st)                      # object API
            assert obj == {"from": st.Get_source(), "world": sub.world_ranks[st.Get_source()]}
    else:
        sub.send({"from": sub.Get_rank(), "world": rank}, dest=0)
world.Barrier()

# 5. Irecv(ANY_SOURCE) + Waitall with statuses; Sendrecv with ANY_SOURCE around a ring
right, left = (rank + 1) % n, (rank - 1) % n
st = MPI.Status()
out = np.empty(4, dtype=np.int64)
world.Sendrecv(np.full(4, rank, dtype=np.int64), dest=right, sendtag=3, recvbuf=out, source=MPI.ANY_SOURCE, recvtag=3, status=st)
assert st.Get_source() == left and np.all(out == left)
world.Barrier()
if rank == 0:
    bufs = [np.empty(2, dtype=np.int64) for _ in range(n - 1)]
    reqs = [world.Irecv(b, source=MPI.ANY_SOURCE, tag=9) for b in bufs]
    sts = [MPI.Status() for _ in reqs]
    MPI.Request.Waitall(reqs, sts)
    assert sorted(s.Get_source() for s in sts) == list(range(1, n))
    for b, s in zip(bufs, sts):
        assert np.all(b == s.Get_source() * 10)
else:
    world.Send(np.full(2, rank * 10, dtype=np.int64), dest=0, tag=9)
world.Barrier()

# 6. prefix reductions and the remaining lowercase forms
assert world.scan(rank + 1) == (rank + 1) * (rank + 2) // 2
assert world.exscan(rank + 1) == (None if rank == 0 else rank * (rank + 1) // 2)
assert world.reduce([rank], root=1) == (list(range(n)) if rank == 1 else None)         # SUM of lists: concatenation, rank order
assert world.allreduce(rank % 2 == 0, MPI.LAND) is False and world.allreduce(rank == 1, MPI.LOR) is True
x = np.arange(6, dtype=np.int64) + rank
inc, exc = np.empty_like(x), np.full_like(x, -7)
world.Scan(x, inc)
world.Exscan(x, exc, op=MPI.MAX)
assert np.array_equal(inc, (rank + 1) * np.arange(6) + rank * (rank + 1) // 2)
assert np.array_equal(exc, np.full(6, -7) if rank == 0 else np.arange(6) + rank - 1)
# object isend / irecv: the receiver learns the size from the header
if rank == 0:
    reqs = [world.irecv(source=j, tag=40 + j) for j in range(1, n)]
    for j, rq in enumerate(reqs, 1):
        st = MPI.Status()
        assert rq.wait(st) == {"rank": j, "payload": "x" * (10 * j)} and st.Get_source() == j
else:
    world.isend({"rank": rank, "payload": "x" * (10 * rank)}, dest=0, tag=40 + rank).Wait()
world.Barrier()
if rank == 0:
    print(f"PROBE STATUS OK ranks={n}")
