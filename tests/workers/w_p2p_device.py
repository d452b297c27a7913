"""Rank-side body of the NVLink point-to-point check (CCMPI_P2P=device): CUDA tensors through
Send/Recv, Sendrecv and Isend/Irecv/Waitall, sizes from a few bytes to 64 MiB, repeated so staging
blocks are reclaimed and reused; plus the reference's hand-rolled collectives on CUDA tensors,
which are built from exactly these calls (mpi_wrapper/comm.py:63-199 in the reference)."""
import numpy as np
import torch

from mpi4py import MPI
from collective_communication_mpi_b200.mpi import p2p_device as P

w = MPI.COMM_WORLD
r, n = w.Get_rank(), w.Get_size()
torch.cuda.set_device(r)
assert P.enabled(), "the NVLink route is the default (CCMPI_P2P=host turns it off)"
nxt, prv = (r + 1) % n, (r - 1) % n

for it, count in enumerate([1, 3, 1000, 1 << 16, (1 << 24) + 5]):
    x = torch.full((count,), float(r * 100 + it), device="cuda")
    y = torch.empty(count, device="cuda")
    # everyone sends, then everyone receives: buffered-send semantics
    w.Send(x, dest=nxt, tag=it)
    x.fill_(-1.0)                                   # the send buffer is ours again
    w.Recv(y, source=prv, tag=it)
    assert bool((y == float(prv * 100 + it)).all()), (r, it, y[:4].tolist())
    z = torch.empty(count, device="cuda", dtype=torch.int32)
    w.Sendrecv(torch.full((count,), r + it, device="cuda", dtype=torch.int32), dest=prv, sendtag=r,
               recvbuf=z, source=nxt, recvtag=nxt)
    assert bool((z == nxt + it).all())

inbox = {j: torch.empty(4096, device="cuda", dtype=torch.int64) for j in range(n) if j != r}
reqs = [w.Irecv(inbox[j], source=j) for j in inbox]
reqs += [w.Isend(torch.arange(4096, device="cuda") + 1000 * r, dest=j) for j in inbox]
MPI.Request.Waitall(reqs)
for j, buf in inbox.items():
    assert torch.equal(buf, torch.arange(4096, device="cuda") + 1000 * j)

led = P.ledger_for(w._world)
w.Barrier()
led.reclaim()
assert led.outstanding == 0, led.outstanding       # every message was acknowledged

# a CUDA receiver also accepts a payload that came the host way (sender had a NumPy array)
if r == 0 and n > 1:
    w.Send(np.arange(100, dtype=np.float32), dest=1, tag=77)
elif r == 1:
    got = torch.empty(100, device="cuda")
    w.Recv(got, source=0, tag=77)
    assert torch.equal(got.cpu(), torch.arange(100, dtype=torch.float32))
w.Barrier()
# ... and a NumPy receiver accepts a CUDA sender's message (descriptor through the ring, payload pulled to the host)
if r == 0 and n > 1:
    w.Send(torch.arange(5000, device="cuda", dtype=torch.int64) * 3, dest=1, tag=78)
elif r == 1:
    host = np.empty(5000, dtype=np.int64)
    w.Recv(host, source=0, tag=78)
    assert np.array_equal(host, np.arange(5000, dtype=np.int64) * 3)
w.Barrier()
led.reclaim()
assert led.outstanding == 0, led.outstanding
if r == 0:
    print(f"P2P DEVICE OK ranks={n}")
