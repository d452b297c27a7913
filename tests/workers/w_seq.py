"""Rank-side body of the sequence-parallel fused-op test: one process per GPU, the stateful
``AllGatherGemmM`` / ``GemmReduceScatterM`` contexts over the world communicator, three launches
each (epochs, double buffering). Oracle: the same math with the library collectives on fp32."""
import torch

from mpi4py import MPI
from collective_communication_mpi_b200.ops.fused_seq import AllGatherGemmM, GemmReduceScatterM

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
torch.cuda.set_device(rank)
BF = torch.bfloat16
rows, K, N = 512, 1024, 768
M = rows * n


def rnd(*shape, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * 0.5).to(BF)


ag = AllGatherGemmM(world, M, K)
rs = GemmReduceScatterM(world, M, N)
w = rnd(K, N, seed=100 + rank)
for it in range(3):
    x = rnd(rows, K, seed=it * 10 + rank)
    y = ag(x, w, out_dtype=torch.float32)
    torch.cuda.synchronize()
    x_all = torch.cat([rnd(rows, K, seed=it * 10 + r) for r in range(n)], dim=0)
    torch.testing.assert_close(y, x_all.float() @ w.float(), rtol=1e-3, atol=2e-2)
    assert torch.equal(ag.gathered(), x_all)

    a = rnd(M, K, seed=500 + it * 10 + rank)
    out = rs(a, w, b_kn=True)
    torch.cuda.synchronize()
    total = sum(rnd(M, K, seed=500 + it * 10 + r).float() @ rnd(K, N, seed=100 + r).float() for r in range(n))
    ref = total[rank * rows:(rank + 1) * rows]
    assert (out.float() - ref).abs().max().item() < 0.15 * max(1.0, ref.abs().max().item() / 10)
    world.Barrier()
if rank == 0:
    print(f"SEQ OK ranks={n}")
