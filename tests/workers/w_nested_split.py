"""Sibling sub-communicators that split again at the same time (8 ranks: world -> halves A, B ->
each half into pairs), then used concurrently with different amounts of work per group.

Checks that every communicator alive in the job has its own id (the barrier slot of the host plane
and the signal pad of the device plane are indexed by it) and that collectives on sibling groups do
not disturb each other. Mirrors the reference's nested use of ``Split`` in ``get_info``
(func_impl.py:57-62), one level deeper.
"""
import sys
import time

import numpy as np

from mpi4py import MPI

world = MPI.COMM_WORLD
rank, size = world.Get_rank(), world.Get_size()
assert size == 8, "run with 8 ranks"
half = world.Split(color=rank // 4, key=rank)             # A = {0..3}, B = {4..7}
pair = half.Split(color=half.Get_rank() // 2, key=rank)   # both halves split at the same moment
dup = half.Dup()
ids = world.allgather((half.comm_id, pair.comm_id, dup.comm_id))
# ids of groups that are different sets of ranks must differ; members of one group agree
by_group = {}
for r, (h, p, d) in enumerate(ids):
    by_group.setdefault(("half", r // 4), set()).add(h)
    by_group.setdefault(("pair", r // 2), set()).add(p)
    by_group.setdefault(("dup", r // 4), set()).add(d)
assert all(len(v) == 1 for v in by_group.values()), by_group
flat = [next(iter(v)) for v in by_group.values()]
assert len(set(flat)) == len(flat) and 0 not in flat, f"communicator ids collide: {by_group}"

# concurrent use: the groups run different numbers of rounds, skewed in time
rounds = 20 + 7 * (rank // 2)
for it in range(rounds):
    if (rank // 2) % 2 == 1 and it % 5 == 0:
        time.sleep(0.002)
    src = np.full(16, rank + it, dtype=np.int64)
    dst = np.empty_like(src)
    pair.Allreduce(src, dst, op=MPI.SUM)
    base = (rank // 2) * 2
    assert (dst == 2 * base + 1 + 2 * it).all(), (rank, it, dst[:2])
    pair.Barrier()
for it in range(10 + 5 * (rank // 4)):
    got = dup.allgather(rank * 100 + it)
    assert got == [r * 100 + it for r in range((rank // 4) * 4, (rank // 4) * 4 + 4)], got
    src = np.arange(8, dtype=np.float64) + rank
    dst = np.empty(2, dtype=np.float64)
    half.Reduce_scatter_block(src, dst, op=MPI.SUM)
    hr = half.Get_rank()
    lo = (rank // 4) * 4
    expect = np.array([sum(j + r for r in range(lo, lo + 4)) for j in (2 * hr, 2 * hr + 1)], dtype=np.float64)
    assert np.array_equal(dst, expect), (rank, dst, expect)
world.Barrier()
if rank == 0:
    print("NESTED SPLIT OK", sorted(flat))
sys.exit(0)
