"""Rank-side check that closing fused ops gives their flag words and symmetric buffers back, and
that a flag region is safe for its next owner whatever the previous one left in it.

Each cycle builds an all-gather->GEMM op and a GEMM->reduce-scatter op with that cycle's token
count, uses both a few times against fp32 math, and closes them. The creation order alternates,
so a region that held counting flags (the gather's ready counters) next holds epoch flags (the
reduce-scatter's slot flags) and the other way round, with a different extent each time."""
import torch

from mpi4py import MPI
from collective_communication_mpi_b200.ops import fused as F

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
torch.cuda.set_device(rank)
DEV = torch.device("cuda", rank)

K, NN = 256 * n, 256 * n
ctx = F._ctx(world)
heap = ctx.heap
live0, used0 = len(heap._live), ctx.flags.used
g = torch.Generator().manual_seed(11)
Wfull = (torch.randn(K, NN, generator=g) * 0.2).bfloat16()


def check(name, got, ref, tol):
    err = (got.float() - ref).abs().max().item()
    lim = tol * max(1.0, ref.abs().max().item())
    assert err < lim, f"{name}: max error {err} > {lim}"


for cycle, M in enumerate([512, 256, 1024, 256, 768, 512]):
    if cycle % 2 == 0:
        ag = F.AllGatherGemm(world, M, K)
        rs = F.GemmReduceScatter(world, M, NN)
    else:
        rs = F.GemmReduceScatter(world, M, NN)
        ag = F.AllGatherGemm(world, M, K)
    if cycle:
        assert min(ag.flag_base, rs.flag_base) == 0, "freed flag words were not handed out again"
    X = (torch.randn(M, K, generator=g) * 0.5).bfloat16()
    w_cols = Wfull[:, rank * (NN // n):(rank + 1) * (NN // n)].contiguous().to(DEV)     # column shard for the gather op
    w_rows = Wfull[rank * (K // n):(rank + 1) * (K // n)].contiguous().to(DEV)          # row shard for the reduce-scatter op
    for it in range(3):
        xs = (X * (it + 1)).bfloat16()
        x_loc = xs[:, rank * (K // n):(rank + 1) * (K // n)].contiguous().to(DEV)
        y = ag(x_loc, w_cols, b_kn=True)
        z = rs(x_loc, w_rows, b_kn=True)
        torch.cuda.synchronize()
        ref = xs.to(DEV).float() @ Wfull.to(DEV).float()
        check(f"cycle {cycle} it {it} ag", y, ref[:, rank * (NN // n):(rank + 1) * (NN // n)], 0.05)
        check(f"cycle {cycle} it {it} rs", z, ref[:, rank * (NN // n):(rank + 1) * (NN // n)], 0.06)
    ag.close()
    rs.close()
    rs.close()                                     # idempotent
    assert ctx.flags.used == used0, "flag words leaked"
    assert len(heap._live) == live0, "symmetric buffers leaked"
    try:
        ag(x_loc, w_cols)
        raise SystemExit("closed op accepted a call")
    except RuntimeError:
        pass

world.Barrier()
if rank == 0:
    print(f"SLOT REUSE OK ranks={n}")
