"""Rank-side check of ``set_tokens``: a block built for one token count is given another one (its fused
ops are closed and rebuilt, weights kept) and still matches the dense fp32 model, forward and backward.
Args: MP [--sp | --megatron]. Runs on CPU ranks (host statements of the ops) and on GPUs (fused kernels;
there the old ops' flag words and symmetric buffers must come back: ops.fused.FusedOp.close)."""
import sys

import torch

from mpi4py import MPI
from collective_communication_mpi_b200.models import TransformerFCBlock, TransformerFCBlockMegatron, TransformerFCBlockSP

mp = int(sys.argv[1])
SP, MEGATRON = "--sp" in sys.argv, "--megatron" in sys.argv
world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
assert n == mp
CUDA = torch.cuda.is_available()
DEV = torch.device("cuda", rank) if CUDA else torch.device("cpu")
if CUDA:
    torch.cuda.set_device(rank)
H = 512
unit = 256 * mp if SP else 256                 # the token-axis kernels want 256 rows per rank
cls = TransformerFCBlockSP if SP else TransformerFCBlockMegatron if MEGATRON else TransformerFCBlock
model = cls(world, unit, H, ffn_mult=4, seed=7)
sh = H // mp


def shard(full, tokens):
    if MEGATRON:
        return full
    rows = tokens // mp
    return full[rank * rows:(rank + 1) * rows] if SP else full[:, rank * sh:(rank + 1) * sh]


def gathered(layer, axis):
    return torch.cat(list(world.allgather(layer.weight.detach().float().cpu())), dim=axis).to(DEV)


Wqkv = [w.to(DEV) for w in world.allgather(model.fc_qkv.weight.detach().float().cpu())]
Wo, W1, W2 = gathered(model.fc_o, 0), gathered(model.fc1, 1), gathered(model.fc2, 0)
W2.requires_grad_(True)


def dense(x):
    a = torch.cat([(lambda q, k, v: v * torch.sigmoid(q * k))(*(x @ w).chunk(3, dim=-1)) for w in Wqkv], dim=-1)
    h = x + a @ Wo
    return h + torch.nn.functional.gelu(h @ W1, approximate="tanh") @ W2


seen_2x = None
if CUDA and mp > 1:
    from collective_communication_mpi_b200.ops.fused import _ctx
    ctx = _ctx(world)
g = torch.Generator().manual_seed(5)
for cycle, tokens in enumerate([unit, 2 * unit, unit, 3 * unit, 2 * unit]):
    model.set_tokens(tokens)                   # first round: same size, nothing to do
    assert all(l.M == tokens for l in model.tp_layers())
    if CUDA and mp > 1 and cycle in (1, 4):    # the same size twice: the second time nothing may have leaked
        now = (len(ctx.heap._live), ctx.flags.used)
        assert cycle == 1 or now == seen_2x, f"buffers or flag words leaked: {seen_2x} -> {now}"
        seen_2x = now
    X = torch.randn(tokens, H, generator=g) * 0.5
    DY = torch.randn(tokens, H, generator=g) * 0.5
    for it in range(2):
        x_loc = shard(X * (it + 1), tokens).contiguous().to(DEV).bfloat16().requires_grad_(True)
        y = model(x_loc)
        for p in model.parameters():
            p.grad = None
        y.backward(shard(DY, tokens).contiguous().to(DEV).bfloat16())
        if CUDA:
            torch.cuda.synchronize()
        x_full = (X * (it + 1)).bfloat16().float().to(DEV)
        W2.grad = None
        ref = dense(x_full)
        ref.backward(DY.bfloat16().float().to(DEV))
        tol = 0.06 * max(1.0, ref.abs().max().item())
        err = (y.float() - shard(ref.detach(), tokens)).abs().max().item()
        assert err < tol, f"tokens={tokens} it={it}: forward error {err} > {tol}"
        dw_ref = W2.grad[rank * (4 * H // mp):(rank + 1) * (4 * H // mp)]
        dw = model.fc2.weight.grad.float()
        errw = (dw - dw_ref).abs().max().item()
        tolw = 0.08 * max(1.0, dw_ref.abs().max().item())
        assert errw < tolw, f"tokens={tokens} it={it}: fc2 dW error {errw} > {tolw}"
world.Barrier()
if rank == 0:
    print(f"SET TOKENS OK mp={mp} sp={int(SP)} megatron={int(MEGATRON)}")
