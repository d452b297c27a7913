"""Stress of the shared-memory transport's flag / ring protocols: ranks arrive at collectives with
random skew, payloads change every iteration, sub-communicators interleave with the world, and
point-to-point traffic of odd sizes crosses ring boundaries."""
from conftest import run_ranks

CODE = r'''
import random, time
import numpy as np
from mpi4py import MPI
w = MPI.COMM_WORLD
r, n = w.Get_rank(), w.Get_size()
rnd = random.Random(1000 + r)
sub = w.Split(color=r % 2, key=r)
for it in range(150):
    if rnd.random() < 0.3:
        time.sleep(rnd.random() * 0.003)                 # skewed arrival
    cnt = 1 + (it * 137) % 9000                      # both sides of the 32 KiB small/large switch
    x = np.full(cnt, r + it, dtype=np.int64); y = np.empty_like(x)
    w.Allreduce(x, y)
    assert y[0] == n * it + n * (n - 1) // 2 and y[-1] == y[0], (it, y[0])
    z = np.empty(cnt, dtype=np.int64)
    sub.Allreduce(x, z, op=MPI.MAX)
    assert z[0] == it + max(q for q in range(n) if q % 2 == r % 2)
    if it % 7 == 0:
        size = 1 + (it * 7919) % 300_000                 # odd byte counts, larger than a ring sometimes
        buf = np.full(size, (r * 31 + it) % 251, dtype=np.uint8); got = np.empty(size, dtype=np.uint8)
        w.Sendrecv(buf, dest=(r + 1) % n, sendtag=it, recvbuf=got, source=(r - 1) % n, recvtag=it)
        assert got[0] == (((r - 1) % n) * 31 + it) % 251 and got[-1] == got[0]
    if it % 5 == 0:                                      # other collectives between the two communicators
        k = sub.Get_size(); sr = sub.Get_rank()
        a2a = np.empty(k * 3, dtype=np.int64)
        sub.Alltoall(np.repeat(np.arange(k, dtype=np.int64) + 100 * sr + it, 3), a2a)
        assert a2a.tolist() == [v for q in range(k) for v in [sr + 100 * q + it] * 3]
        ag = np.empty(n * 5, dtype=np.float64)
        w.Allgather(np.full(5, r * 2.0 + it), ag)
        assert ag.tolist() == [q * 2.0 + it for q in range(n) for _ in range(5)]
        rs = np.empty(4, dtype=np.int64)
        w.Reduce_scatter_block(np.arange(4 * n, dtype=np.int64) + r, rs)
        assert rs.tolist() == [n * (4 * r + j) + n * (n - 1) // 2 for j in range(4)]
    g = w.allgather((r, it))
    assert g == [(q, it) for q in range(n)]
w.Barrier()
if r == 0: print("STRESS OK", n)
'''


def test_skewed_arrivals_changing_payloads_and_subgroups():
    p = run_ranks(6, ["-c", CODE], timeout=240)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "STRESS OK 6" in p.stdout
