"""Rank-side check of ``ColumnParallelLinear(gather_output=True)`` -- the reference's complete
forward step, gather the input / multiply / gather the output (func_impl.py:76-109), and its
backward (slice of the output gradient, func_impl.py:111-147) -- against dense fp32 math.
Runs the default composition, or the single-kernel flavour with CCMPI_FUSED_GATHER_OUTPUT=1."""
import os

import torch

from mpi4py import MPI
from collective_communication_mpi_b200.parallel.tp_layers import ColumnParallelLinear

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
CUDA = torch.cuda.is_available()
DEV = torch.device("cuda", rank) if CUDA else torch.device("cpu")
if CUDA:
    torch.cuda.set_device(rank)


def sync():
    if CUDA:
        torch.cuda.synchronize()


TOK, K, NOUT = 512, 256 * n, 128 * n
layer = ColumnParallelLinear(world, TOK, K, NOUT, gather_output=True, seed=3)
g = torch.Generator().manual_seed(5)
X = (torch.randn(TOK, K, generator=g) * 0.5).bfloat16()
DY = (torch.randn(TOK, NOUT, generator=g) * 0.5).bfloat16()
sh = K // n
W = torch.cat(list(world.allgather(layer.weight.detach().float().cpu())), dim=1).to(DEV)      # [K, NOUT]
for it in range(3):
    x_loc = (X[:, rank * sh:(rank + 1) * sh] * (it + 1)).contiguous().to(DEV).requires_grad_(True)
    y = layer(x_loc)
    sync()
    x_full = (X * (it + 1)).bfloat16().to(DEV).float()
    ref = x_full @ W
    assert y.shape == (TOK, NOUT)
    assert (y.float() - ref).abs().max().item() < 0.06 * max(1.0, ref.abs().max().item()), "forward"
    layer.weight.grad = None
    y.backward(DY.to(DEV))
    sync()
    dy = DY.to(DEV).float()
    dx_ref = (dy @ W.t())[:, rank * sh:(rank + 1) * sh]          # every rank contributes its column block of dy
    dw_ref = x_full.t() @ dy[:, rank * (NOUT // n):(rank + 1) * (NOUT // n)]
    assert (x_loc.grad.float() - dx_ref).abs().max().item() < 0.08 * max(1.0, dx_ref.abs().max().item()), "dx"
    assert (layer.weight.grad.float() - dw_ref).abs().max().item() < 0.08 * max(1.0, dw_ref.abs().max().item()), "dW"
world.Barrier()
if rank == 0:
    print(f"GATHER OUTPUT OK ranks={n} fused={os.environ.get('CCMPI_FUSED_GATHER_OUTPUT', '0')}")
