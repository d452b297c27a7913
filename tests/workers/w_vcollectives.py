"""The "v" collectives (blocks of different sizes) of the mpi4py front end, NumPy and -- with ``--torch`` --
CPU torch tensors, against values every rank can compute by itself."""
import sys

import numpy as np

from mpi4py import MPI

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
TORCH = "--torch" in sys.argv
if TORCH:
    import torch


def mk(a):
    return torch.from_numpy(np.ascontiguousarray(a)) if TORCH else a


def host(a):
    return a.numpy() if TORCH else a


counts = [(3 * j) % 5 + (1 if j != 2 else 0) for j in range(n)]          # ragged, rank 2 sends nothing
total = sum(counts)
block = lambda j: np.arange(counts[j], dtype=np.float64) + 100.0 * j     # noqa: E731
full = np.concatenate([block(j) for j in range(n)])

# Allgatherv: packed, and with gaps (displacements)
out = mk(np.full(total, -1.0))
world.Allgatherv(mk(block(rank)), [out, counts])
assert np.array_equal(host(out), full)
displs = [2 * sum(counts[:j]) + 1 for j in range(n)]
out = mk(np.full(2 * total + 2, -1.0))
world.Allgatherv(mk(block(rank)), [out, counts, displs, MPI.DOUBLE])
for j in range(n):
    assert np.array_equal(host(out)[displs[j]:displs[j] + counts[j]], block(j))
assert host(out)[0] == -1.0 and np.sum(host(out) == -1.0) == out.shape[0] - total
# plain buffers: equal blocks, like Allgather
eq = mk(np.empty(3 * n, dtype=np.int64))
world.Allgatherv(mk(np.full(3, rank, dtype=np.int64)), eq)
assert np.array_equal(host(eq), np.repeat(np.arange(n), 3))

# Gatherv: only the root holds counts and a receive buffer
root = n - 1
if rank == root:
    got = mk(np.full(total, -1.0))
    world.Gatherv(mk(block(rank)), [got, counts], root=root)
    assert np.array_equal(host(got), full)
else:
    world.Gatherv(mk(block(rank)), None, root=root)

# Scatterv: the inverse
mine = mk(np.full(max(counts) + 1, -1.0))
if rank == 0:
    world.Scatterv([mk(full), counts], mine, root=0)
else:
    world.Scatterv(None, mine, root=0)
assert np.array_equal(host(mine)[:counts[rank]], block(rank)) and host(mine)[counts[rank]] == -1.0

# Alltoallv: rank i sends (i + j) % 3 elements to rank j
sc = [(rank + j) % 3 for j in range(n)]
rc = [(i + rank) % 3 for i in range(n)]
send = np.concatenate([np.full(sc[j], 10 * rank + j, dtype=np.int32) for j in range(n)] + [np.empty(0, dtype=np.int32)])
recv = mk(np.full(sum(rc), -1, dtype=np.int32))
world.Alltoallv([mk(send), sc], [recv, rc])
want = np.concatenate([np.full(rc[i], 10 * i + rank, dtype=np.int32) for i in range(n)] + [np.empty(0, dtype=np.int32)])
assert np.array_equal(host(recv), want)

# Reduce_scatter with unequal recvcounts
vec = np.arange(total, dtype=np.int64) * (rank + 1)
seg = mk(np.empty(counts[rank], dtype=np.int64))
world.Reduce_scatter(mk(vec), seg, recvcounts=counts)
lo = sum(counts[:rank])
assert np.array_equal(host(seg), np.arange(total)[lo:lo + counts[rank]] * (n * (n + 1) // 2))

# MPI.IN_PLACE: the contribution already sits in the receive buffer
buf = np.full(3 * n, -1, dtype=np.int64)
buf[3 * rank:3 * rank + 3] = rank
bufm = mk(buf)
world.Allgather(MPI.IN_PLACE, bufm)
assert np.array_equal(host(bufm), np.repeat(np.arange(n), 3))
acc = mk(np.arange(5, dtype=np.float64) + rank)
world.Allreduce(MPI.IN_PLACE, acc)
assert np.allclose(host(acc), n * np.arange(5) + n * (n - 1) / 2)

# errors are raised on every rank before any traffic
try:
    world.Allgatherv(mk(block(rank)), [mk(np.empty(total)), counts[:-1]])
    raise SystemExit("short counts accepted")
except ValueError:
    pass
world.Barrier()
if rank == 0:
    print(f"V COLLECTIVES OK ranks={n} torch={int(TORCH)}")
