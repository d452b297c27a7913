"""Train the transformer-FC block on a 2-D (data-parallel x model-parallel) grid.

    # 4 ranks = 2 replicas x 2 model-parallel shards; with GPUs: one per rank, fused kernels over NVLink;
    # without: CPU tensors over the shared-memory transport (same code, ops/host_ops.py underneath)
    python -m collective_communication_mpi_b200.launch -n 4 python examples/train_block.py --mp 2 --steps 5

Uses only the public API a user of the reference already knows -- ``MPI.COMM_WORLD``, ``get_info`` (the
reference's model-parallel-major grid, model/func_impl.py:5-74) and ``split_data``
(data/data_parallel_preprocess.py:3-59) -- plus this framework's layers and ``TrainStep``.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mpi4py import MPI  # noqa: E402
from data.data_parallel_preprocess import split_data  # noqa: E402
from model.func_impl import get_info  # noqa: E402
from collective_communication_mpi_b200.models import (TrainStep, TransformerFCBlock, TransformerFCBlockMegatron,  # noqa: E402
                                                      TransformerFCBlockSP, TransformerFCStack)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mp", type=int, default=2, help="model-parallel size (ranks per replica)")
    ap.add_argument("--hidden", type=int, default=512)
    ap.add_argument("--tokens", type=int, default=256, help="tokens per replica and step")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--depth", type=int, default=1, help="number of blocks in sequence (reference layout; the stack shares one "
                                                         "reduce-scatter op between all its layers)")
    ap.add_argument("--sp", action="store_true", help="sequence-parallel layout: each rank holds tokens/mp rows of the full width")
    ap.add_argument("--megatron", action="store_true", help="classic Megatron layout: activations replicated across the mp group")
    ap.add_argument("--lr", type=float, default=None, help="plain SGD on a mean-squared error: gradients are tiny "
                                                           "(default 10 for one block, 0.3 for a stack)")
    args = ap.parse_args()
    if args.lr is None:
        args.lr = 10.0 if args.depth == 1 else 0.3

    comm = MPI.COMM_WORLD
    rank, world = comm.Get_rank(), comm.Get_size()
    if world % args.mp:
        raise SystemExit(f"--mp {args.mp} does not divide the {world} ranks")
    dp = world // args.mp
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(rank % torch.cuda.device_count())
    dev = torch.device("cuda") if cuda else torch.device("cpu")

    mp_idx, dp_idx, mp_comm, dp_comm, _, _ = get_info(comm=comm, rank=rank, mp_size=args.mp, dp_size=dp, fc_layer="fc_q",
                                                      in_dim=args.hidden, out_dim=args.hidden)
    block = TransformerFCBlockSP if args.sp else TransformerFCBlockMegatron if args.megatron else TransformerFCBlock
    if args.depth > 1:
        if args.sp or args.megatron:
            raise SystemExit("--depth > 1 builds the reference-layout stack; drop --sp / --megatron")
        model = TransformerFCStack(mp_comm, args.tokens, args.hidden, args.depth, ffn_mult=4, seed=1)
    else:
        model = block(mp_comm, args.tokens, args.hidden, ffn_mult=4, seed=1)
    step = TrainStep(model, dp_comm, lr=args.lr)

    # one fixed synthetic regression task; every replica gets its slice of the batch, every model-parallel
    # rank its slice of the features
    g = torch.Generator().manual_seed(0)
    X = torch.randn(dp * args.tokens, args.hidden, generator=g) * 0.5
    Y = torch.tanh(X.roll(1, dims=1)) * 0.5
    xs, ys = split_data(X, Y, mp_size=args.mp, dp_size=dp, rank=rank)
    if args.megatron:  # every rank of the replica holds the whole batch
        x, y = xs.contiguous().to(dev).bfloat16(), ys.contiguous().to(dev).bfloat16()
    elif args.sp:     # this rank's block of rows, all features
        rows = args.tokens // args.mp
        x = xs[mp_idx * rows:(mp_idx + 1) * rows].contiguous().to(dev).bfloat16()
        y = ys[mp_idx * rows:(mp_idx + 1) * rows].contiguous().to(dev).bfloat16()
    else:             # the reference's layout: all rows, this rank's block of features
        sh = args.hidden // args.mp
        x = xs[:, mp_idx * sh:(mp_idx + 1) * sh].contiguous().to(dev).bfloat16()
        y = ys[:, mp_idx * sh:(mp_idx + 1) * sh].contiguous().to(dev).bfloat16()

    losses = []
    for _ in range(args.steps):
        local = float(step(x, y))
        losses.append(comm.allreduce(local, op=MPI.SUM) / world)      # mean over shards and replicas
    if rank == 0:
        print(f"grid mp{args.mp} x dp{dp}{' (sequence-parallel)' if args.sp else ' (Megatron layout)' if args.megatron else ''} on {'GPU' if cuda else 'CPU'}: loss " + " -> ".join(f"{v:.4f}" for v in losses))
        print("TRAIN OK" if losses[-1] < losses[0] else "TRAIN DID NOT IMPROVE")
    comm.Barrier()


if __name__ == "__main__":
    main()
