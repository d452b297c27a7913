"""Rank-side check of the tensor-parallel layers / model / train step against a dense fp32
PyTorch model assembled from the gathered weight shards. Args: MP DP [--sp].
``--sp``: the sequence-parallel block (activations sharded by tokens outside the GEMM pairs)."""
import sys

import torch

from mpi4py import MPI
from model.func_impl import get_info
from data.data_parallel_preprocess import split_data
from collective_communication_mpi_b200.models import (TrainStep, TransformerFCBlock, TransformerFCBlockMegatron,
                                                      TransformerFCBlockSP)

mp, dp = int(sys.argv[1]), int(sys.argv[2])
SP = "--sp" in sys.argv
MEGATRON = "--megatron" in sys.argv        # replicated activations, GEMM -> all-reduce
world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
assert n == mp * dp
CUDA = torch.cuda.is_available()
DEV = torch.device("cuda", rank) if CUDA else torch.device("cpu")
if CUDA:
    torch.cuda.set_device(rank)


def sync():
    if CUDA:
        torch.cuda.synchronize()


mp_idx, dp_idx, mp_comm, dp_comm, _, _ = get_info(comm=world, rank=rank, mp_size=mp, dp_size=dp, fc_layer="fc_q",
                                                  in_dim=1024, out_dim=1024)
H, TOK = (512 if mp <= 4 else 1024), 256     # every scattered chunk must be >= 128 wide
if SP:
    TOK = 256 * mp                            # the kernels want 256 rows per rank
    model = TransformerFCBlockSP(mp_comm, TOK, H, ffn_mult=4, seed=7)
elif MEGATRON:
    model = TransformerFCBlockMegatron(mp_comm, TOK, H, ffn_mult=4, seed=7)
else:
    model = TransformerFCBlock(mp_comm, TOK, H, ffn_mult=4, seed=7, num_comm_ctas=2)

# global batch: dp replicas x TOK tokens, feature-sharded over mp
g = torch.Generator().manual_seed(123)
X = torch.randn(dp * TOK, H, generator=g) * 0.5
T = torch.randn(dp * TOK, H, generator=g) * 0.5
xs, ts = split_data(X, T, mp_size=mp, dp_size=dp, rank=rank)
sh = H // mp
rows = TOK // mp


def shard(full, r):
    """Rank r's part of a [TOK, H] activation: a column block, or a row block with --sp."""
    if MEGATRON:
        return full                           # replicated: every rank holds the whole activation
    return full[r * rows:(r + 1) * rows] if SP else full[:, r * sh:(r + 1) * sh]


x_loc = shard(xs, mp_idx).contiguous().to(DEV).bfloat16()
t_loc = shard(ts, mp_idx).contiguous().to(DEV).bfloat16()


def full_weight(layer, axis):
    parts = mp_comm.allgather(layer.weight.detach().float().cpu())
    return torch.cat(list(parts), dim=axis).to(DEV)


Wqkv_parts = mp_comm.allgather(model.fc_qkv.weight.detach().float().cpu())   # each [H, 3H/mp] = q|k|v shards
Wo, W1, W2 = full_weight(model.fc_o, 0), full_weight(model.fc1, 1), full_weight(model.fc2, 0)


def dense_forward(x_full):
    """Same math on unsharded fp32 tensors; q/k/v columns are interleaved per rank like the shards."""
    outs = []
    for r in range(mp):
        q, k, v = (x_full @ Wqkv_parts[r].to(DEV)).chunk(3, dim=-1)
        outs.append(v * torch.sigmoid(q * k))
    a = torch.cat(outs, dim=-1)
    h = x_full + a @ Wo
    u = torch.nn.functional.gelu(h @ W1, approximate="tanh")
    return h + u @ W2


x_full = xs.to(DEV).float().bfloat16().float()
ref = dense_forward(x_full)
y = model(x_loc)
sync()
got = y.float()
want = shard(ref, mp_idx)
err = (got - want).abs().max().item()
assert err < 0.06 * max(1.0, want.abs().max().item()), f"forward mismatch {err}"

# ---- one training step: gradients (averaged over dp) vs autograd on the dense model ------------
before = [l.weight.detach().clone() for l in model.tp_layers()]
LR = 64.0       # large on purpose: the update must be many bf16 ulps of the weights to be measurable
step = TrainStep(model, dp_comm, lr=LR, bucket_mb=1)
loss = step(x_loc, t_loc)
sync()

params = [p.clone().requires_grad_(True) for p in [*[w.to(DEV) for w in Wqkv_parts], Wo, W1, W2]]


def dense_loss(x_full, t_full):
    outs = []
    for r in range(mp):
        q, k, v = (x_full @ params[r]).chunk(3, dim=-1)
        outs.append(v * torch.sigmoid(q * k))
    a = torch.cat(outs, dim=-1)
    h = x_full + a @ params[mp]
    u = torch.nn.functional.gelu(h @ params[mp + 1], approximate="tanh")
    yy = h + u @ params[mp + 2]
    # every rank's loss is the mean over ITS shard; the model-parallel sum of those is mp x the
    # mean over the full width
    if MEGATRON:      # replicated output: every rank evaluates the one and only loss
        return torch.nn.functional.mse_loss(yy, t_full)
    return sum(torch.nn.functional.mse_loss(shard(yy, r), shard(t_full, r)) for r in range(mp))


total = 0.0
for d in range(dp):
    xd = X[d * TOK:(d + 1) * TOK].to(DEV).bfloat16().float()
    td = T[d * TOK:(d + 1) * TOK].to(DEV).bfloat16().float()
    total = total + dense_loss(xd, td) / dp
total.backward()
g_qkv = params[mp_idx].grad
g_o = params[mp].grad[mp_idx * sh:(mp_idx + 1) * sh]
g_1 = params[mp + 1].grad[:, mp_idx * (4 * H // mp):(mp_idx + 1) * (4 * H // mp)]
g_2 = params[mp + 2].grad[mp_idx * (4 * H // mp):(mp_idx + 1) * (4 * H // mp)]
# (a sink whose reduction is the optimizer step never materialises the averaged gradient: only the weights can be checked)
sink_grads = step.sink.grads() if step.sink is not None and not step.sink.fused_update else [None] * 4
for name, layer, w0, gref, gbuf in zip(("qkv", "o", "fc1", "fc2"), model.tp_layers(), before, (g_qkv, g_o, g_1, g_2), sink_grads):
    scale = gref.abs().max().item()
    if gbuf is not None:
        # the reduced gradient itself (bf16, averaged over dp): a few parts in a thousand of its scale
        e = (gbuf.float() - gref).abs().max().item()
        if rank == 0 and "--verbose" in sys.argv:
            print(f"  {name}: max|grad| {scale:.3e}  max|grad err| {e:.3e}")
        assert e < 0.05 * scale, f"{name}: gradient mismatch {e} (scale {scale})"
    # what the optimizer applied: w0 - LR * grad, rounded to bf16
    want = w0.float() - LR * gref
    e = (layer.weight.detach().float() - want).abs().max().item()
    ulp = want.abs().max().item() * 2.0 ** -8
    if rank == 0 and "--verbose" in sys.argv:
        print(f"  {name}: max|update| {LR * scale:.3e}  max|weight err| {e:.3e}  (bf16 ulp {ulp:.1e})")
    assert e < 0.05 * LR * scale + 2 * ulp, f"{name}: weight update mismatch {e} (update scale {LR * scale}, ulp {ulp})"
world.Barrier()
if rank == 0:
    print(f"LAYERS OK mp={mp} dp={dp} sp={int(SP)} megatron={int(MEGATRON)} loss={loss.item():.4f} ar_launches={step.sink.launches if step.sink else 0} fused_sgd={int(step.fused_sgd)} fused_update={int(bool(step.sink and step.sink.fused_update))}")
