"""Rank-side check of ``TransformerFCStack`` (several blocks, ONE shared GEMM->reduce-scatter op for the whole
stack) against the dense fp32 model built from the gathered shards, and of a few training steps. Args: MP DP DEPTH."""
import sys

import torch

from mpi4py import MPI
from model.func_impl import get_info
from collective_communication_mpi_b200.models import TrainStep, TransformerFCStack

mp, dp, depth = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
assert n == mp * dp
CUDA = torch.cuda.is_available()
DEV = torch.device("cuda", rank) if CUDA else torch.device("cpu")
if CUDA:
    torch.cuda.set_device(rank)
mp_idx, dp_idx, mp_comm, dp_comm, _, _ = get_info(comm=world, rank=rank, mp_size=mp, dp_size=dp, fc_layer="fc_q",
                                                  in_dim=512, out_dim=512)
H, TOK = 512, 256
model = TransformerFCStack(mp_comm, TOK, H, depth, ffn_mult=4, seed=3)
layers = model.tp_layers()
assert len(layers) == 4 * depth
if mp > 1:
    scatter_ops = {id(l.rs) for l in layers}
    gather_ops = {id(l.ag) for l in layers}
    assert len(scatter_ops) == 1, "every reduce-scatter of the stack produces [tokens, hidden]: one shared op"
    assert len(gather_ops) == 4 * depth, "gather ops keep their gathered operand for backward: never shared"
    assert layers[0].rs._refs == 4 * depth
    if CUDA:
        from collective_communication_mpi_b200.ops.fused import _ctx
        used = _ctx(mp_comm).flags.used
        per_gather = max(16, (mp * ((TOK + 127) // 128) + 15) // 16 * 16)
        assert used <= 4 * depth * per_gather + 64 * mp + 64, f"{used} flag words in use: scatter regions are not shared"
sh = H // mp


def cat(layer, axis):
    return torch.cat(list(mp_comm.allgather(layer.weight.detach().float().cpu())), dim=axis).to(DEV)


dense = []
for b in model.blocks:
    dense.append(([w.to(DEV) for w in mp_comm.allgather(b.fc_qkv.weight.detach().float().cpu())], cat(b.fc_o, 0),
                  cat(b.fc1, 1), cat(b.fc2, 0).requires_grad_(True)))


def dense_forward(x):
    for wqkv, wo, w1, w2 in dense:
        a = torch.cat([(lambda q, k, v: v * torch.sigmoid(q * k))(*(x @ w).chunk(3, dim=-1)) for w in wqkv], dim=-1)
        h = x + a @ wo
        x = h + torch.nn.functional.gelu(h @ w1, approximate="tanh") @ w2
    return x


g = torch.Generator().manual_seed(50 + dp_idx)
X = torch.randn(TOK, H, generator=g) * 0.5
DY = torch.randn(TOK, H, generator=g) * 0.1
x_loc = X[:, mp_idx * sh:(mp_idx + 1) * sh].contiguous().to(DEV).bfloat16().requires_grad_(True)
y = model(x_loc)
y.backward(DY[:, mp_idx * sh:(mp_idx + 1) * sh].contiguous().to(DEV).bfloat16())
if CUDA:
    torch.cuda.synchronize()
ref = dense_forward(X.bfloat16().float().to(DEV))
ref.backward(DY.bfloat16().float().to(DEV))
tol = 0.03 * depth * max(1.0, ref.abs().max().item())
err = (y.float() - ref.detach()[:, mp_idx * sh:(mp_idx + 1) * sh]).abs().max().item()
assert err < tol, f"forward through {depth} blocks: error {err} > {tol}"
for i in (0, depth - 1):
    dw_ref = dense[i][3].grad[mp_idx * (4 * H // mp):(mp_idx + 1) * (4 * H // mp)]
    dw = model.blocks[i].fc2.weight.grad.float()
    errw = (dw - dw_ref).abs().max().item()
    tolw = 0.05 * depth * max(1.0, dw_ref.abs().max().item())
    assert errw < tolw, f"block {i} fc2 dW: error {errw} > {tolw}"

# a few optimisation steps through the 2-D grid
for p in model.parameters():
    p.grad = None
step = TrainStep(model, dp_comm, lr=2e-2)
xb = x_loc.detach()
tb = (torch.randn(TOK, sh, generator=g) * 0.5).to(DEV).bfloat16()
losses = [float(step(xb, tb)) for _ in range(6)]
assert all(l == l for l in losses) and losses[-1] < losses[0], losses
# closing: every holder closes once, the shared op goes with the last one
from collective_communication_mpi_b200.ops.fused import close_fused_ops     # noqa: E402
if mp > 1:
    shared = layers[0].rs
    assert close_fused_ops(model) == 8 * depth            # one close per holder: 4 gathers + 4 x the shared scatter per block
    assert shared._closed and shared._refs == 1
world.Barrier()
if rank == 0:
    print(f"STACK OK mp={mp} dp={dp} depth={depth} loss {losses[0]:.4f} -> {losses[-1]:.4f}")
