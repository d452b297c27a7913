"""A/B timing of the 1-GPU training step under different switches, robust to the clock drift of a
power-capped GPU: the variants are built side by side in ONE process and timed in interleaved rounds
(A B C A B C ...), each round after the same idle gap, and the per-variant median over rounds is
reported together with every round's number. Not a benchmark of record -- a tool for deciding defaults.

    python benchmarks/ab_step.py "CCMPI_FUSED_EPILOGUE=1" "CCMPI_FUSED_EPILOGUE=0" [--rounds 6 --steps 10 --idle 1.0]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("CCMPI_TRANSPORT", "device")
os.environ.setdefault("CCMPI_HEAP_MB", "4096")
import torch  # noqa: E402
from mpi4py import MPI  # noqa: E402
from model.func_impl import get_info  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("variants", nargs="+", help='each "K=V,K2=V2" (environment for that variant) or "nccl"')
    ap.add_argument("--rounds", type=int, default=6)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--idle", type=float, default=1.0)
    args = ap.parse_args()
    comm = MPI.COMM_WORLD
    _, _, mpc, dpc, _, _ = get_info(comm=comm, rank=0, mp_size=1, dp_size=1, fc_layer="fc_q", in_dim=4096, out_dim=4096)
    TOK, H = 8192, 4096
    from collective_communication_mpi_b200.models import TrainStep, TransformerFCBlock
    steps = []
    for v in args.variants:
        if v == "nccl":
            sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
            from nccl_baseline import NcclTrainStep
            steps.append(NcclTrainStep(0, 1, 1, 1, TOK, H, 4))
            continue
        env = dict(kv.split("=", 1) for kv in v.split(",") if kv)
        old = {k: os.environ.get(k) for k in env}
        os.environ.update(env)
        steps.append(TrainStep(TransformerFCBlock(mpc, TOK, H, 4, seed=1), dpc, lr=1e-4))
        for k, o in old.items():
            if o is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = o
    x = [(torch.randn(TOK, H, device="cuda") * 0.5).bfloat16() for _ in range(4)]
    t = [(torch.randn(TOK, H, device="cuda") * 0.5).bfloat16() for _ in range(4)]
    res = [[] for _ in steps]
    for r in range(args.rounds):
        order = list(range(len(steps)))
        if r % 2:
            order.reverse()         # no variant always runs first (coolest) in a round
        for i in order:
            torch.cuda.synchronize()
            time.sleep(args.idle)
            for k in range(3):
                steps[i](x[k % 4], t[k % 4])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for k in range(args.steps):
                steps[i](x[k % 4], t[k % 4])
            e1.record()
            torch.cuda.synchronize()
            res[i].append(round(e0.elapsed_time(e1) / args.steps, 4))
    for v, rs in zip(args.variants, res):
        srt = sorted(rs)
        print(json.dumps({"variant": v, "median_ms": srt[len(srt) // 2], "min_ms": srt[0], "rounds_ms": rs}), flush=True)


if __name__ == "__main__":
    main()
