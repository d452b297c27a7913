"""BASELINE.md configurations 3 and 4 (and any other grid), ours vs NCCL+cuBLAS.

  tp8   mp=8 tensor-parallel transformer FC block fwd+bwd, hidden 8192, FFN 4x, 8192 tokens, bf16
        (config 3 without its fp8 part: the block-scaled fp8 GEMM is not built yet)
  dp8   dp=N data-parallel MLP stack, 16 x 8192^2 = 1.07 B parameters, 4096 tokens per replica;
        gradient all-reduce fused with the 1/dp scale and overlapped with backward (config 4)
  tp24  the bench.py configuration (mp=2 x dp=N/2, hidden 4096)

    python -m torch.distributed.run --nproc-per-node 8 ... benchmarks/config_bench.py --config tp8 --impl ours

One JSON line on rank 0; device-timed (CUDA events), barrier + synchronize both sides, max over ranks.
"""
import argparse
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
import torch  # noqa: E402

CONFIGS = {
    "tp8": dict(model="block", hidden=8192, ffn=4, tokens=8192, mp="all"),
    "tp24": dict(model="block", hidden=4096, ffn=4, tokens=8192, mp=2),
    "dp8": dict(model="mlp", hidden=8192, depth=16, tokens=4096, mp=1),
    # the same block in the other two activation layouts (ours only)
    "tp8sp": dict(model="block_sp", hidden=8192, ffn=4, tokens=8192, mp="all"),
    "tp8meg": dict(model="block_megatron", hidden=8192, ffn=4, tokens=8192, mp="all"),
    "tp24sp": dict(model="block_sp", hidden=4096, ffn=4, tokens=8192, mp=2),
    "tp24meg": dict(model="block_megatron", hidden=4096, ffn=4, tokens=8192, mp=2),
}


class NcclMLPStep:
    """DDP-style baseline for the MLP stack: torch.matmul + async NCCL all-reduce per layer gradient."""

    def __init__(self, rank, world, width, depth, lr):
        import torch.distributed as dist
        self.dist, self.dp, self.lr = dist, world, lr
        dev = torch.device("cuda", torch.cuda.current_device())
        g = torch.Generator(device=dev).manual_seed(0)
        self.ws = [(torch.randn(width, width, device=dev, generator=g) / math.sqrt(width)).bfloat16().requires_grad_(True)
                   for _ in range(depth)]
        self.pending = []
        if world > 1:
            for w in self.ws:
                w.register_post_accumulate_grad_hook(lambda p: self.pending.append(dist.all_reduce(p.grad, async_op=True)))

    def __call__(self, x, t):
        for w in self.ws:
            w.grad = None
        h = x
        for i, w in enumerate(self.ws):
            h = h @ w
            if i + 1 < len(self.ws):
                h = torch.relu(h)
        loss = torch.nn.functional.mse_loss(h.float(), t.float())
        loss.backward()
        for work in self.pending:
            work.wait()
        self.pending.clear()
        with torch.no_grad():
            grads = [w.grad for w in self.ws]
            if self.dp > 1:
                torch._foreach_mul_(grads, 1.0 / self.dp)
            torch._foreach_add_([w.data for w in self.ws], grads, alpha=-self.lr)
        return loss.detach()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="tp8", choices=list(CONFIGS))
    ap.add_argument("--impl", default="ours", choices=["ours", "nccl"])
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", str(rank))))
    os.environ.setdefault("CCMPI_HEAP_MB", "16384")
    os.environ.setdefault("CCMPI_TRANSPORT", "device")
    from mpi4py import MPI
    from model.func_impl import get_info
    from collective_communication_mpi_b200 import _native as N
    comm = MPI.COMM_WORLD
    mp = world if cfg["mp"] == "all" else min(cfg["mp"], world)
    dp = world // mp
    H, tokens = cfg["hidden"], cfg["tokens"]
    mp_idx, dp_idx, mp_comm, dp_comm, _, _ = get_info(comm=comm, rank=rank, mp_size=mp, dp_size=dp, fc_layer="fc_q",
                                                      in_dim=H, out_dim=H)
    if args.impl == "ours":
        from collective_communication_mpi_b200 import models as Mo
        blocks = {"block": Mo.TransformerFCBlock, "block_sp": Mo.TransformerFCBlockSP, "block_megatron": Mo.TransformerFCBlockMegatron}
        model = blocks[cfg["model"]](mp_comm, tokens, H, cfg["ffn"], seed=1) if cfg["model"] in blocks else \
            Mo.MLPStack(mp_comm, tokens, H, cfg["depth"], seed=1)
        step = Mo.TrainStep(model, dp_comm, lr=1e-4)
    else:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29541")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", torch.cuda.current_device()))
        if cfg["model"] != "mlp" and cfg["model"] != "block":
            raise SystemExit(f"config {args.config} has no NCCL arm: compare with tp8 / tp24 --impl nccl")
        if cfg["model"] == "block":
            from nccl_baseline import NcclTrainStep
            step = NcclTrainStep(rank, world, mp, dp, tokens, H, cfg["ffn"], lr=1e-4)
        else:
            step = NcclMLPStep(rank, world, H, cfg["depth"], 1e-4)
    # input shard per rank: features (reference layout), rows (sequence parallel) or everything (Megatron)
    shape = {"block_sp": (tokens // mp, H), "block_megatron": (tokens, H)}.get(cfg["model"], (tokens, H // mp))
    xs = [(torch.randn(*shape, device="cuda") * 0.5).bfloat16() for _ in range(2)]
    ts = [(torch.randn(*shape, device="cuda") * 0.5).bfloat16() for _ in range(2)]
    for i in range(max(3, args.warmup)):
        step(xs[i & 1], ts[i & 1])
    torch.cuda.synchronize(); comm.Barrier(); torch.cuda.synchronize()
    before = N.LAUNCHES[0]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(xs[i & 1], ts[i & 1])
    e1.record()
    torch.cuda.synchronize(); comm.Barrier()
    ms = max(comm.allgather(e0.elapsed_time(e1))) / args.steps
    if cfg["model"].startswith("block"):
        flops = 3 * 2 * tokens * (H * 3 * H + H * H + 2 * H * cfg["ffn"] * H) * dp
        params = (H * 3 * H + H * H + 2 * H * cfg["ffn"] * H)
    else:
        flops = 3 * 2 * tokens * H * H * cfg["depth"] * dp
        params = H * H * cfg["depth"]
    if rank == 0:
        print(json.dumps({"config": args.config, "impl": args.impl, "n_gpus": world, "parallelism": f"mp{mp}xdp{dp}",
                          "ms_per_step": round(ms, 4), "tokens_per_s": round(tokens * dp / (ms * 1e-3), 1),
                          "gemm_tflops_per_gpu": round(flops / (ms * 1e-3) / 1e12 / world, 1), "params": params,
                          "hidden": H, "tokens_per_replica": tokens, "dtype": "bf16", "steps": args.steps,
                          "gpu_launches": (N.LAUNCHES[0] - before) if args.impl == "ours" else 0}), flush=True)
    comm.Barrier()
    if args.impl == "nccl":
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
