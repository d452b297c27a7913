"""Fused GEMM+collective ops in isolation vs their unfused baselines, per FC shape.

For every (op, shape): our fused kernel | our plain GEMM alone | cuBLAS alone | NCCL collective alone |
NCCL + cuBLAS back to back (incl. the layout copies the last-axis collectives need). Device-timed
(CUDA events), max over ranks, L2 flushed between iterations. Roofline fraction =
max(flops / measured cuBLAS peak, NVLink bytes / 770 GB/s) / fused time.

    python -m collective_communication_mpi_b200.launch -n 2 python benchmarks/fused_bench.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

os.environ.setdefault("CCMPI_HEAP_MB", "8192")
os.environ.setdefault("CCMPI_TRANSPORT", "device")
from mpi4py import MPI  # noqa: E402
from collective_communication_mpi_b200.ops import fused as F  # noqa: E402
from collective_communication_mpi_b200.ops.gemm import gemm  # noqa: E402
from collective_communication_mpi_b200.utils import ktrace as KT  # noqa: E402

PEAK_TFLOPS = 1670.0   # MEASURED_PEAKS.json bf16 burst
LINK_GBS = 770.0       # measured peer copy, per direction


def main():
    comm = MPI.COMM_WORLD
    rank, n = comm.Get_rank(), comm.Get_size()
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    H = int(os.environ.get("FB_HIDDEN", "4096"))
    M = int(os.environ.get("FB_TOKENS", "8192"))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timeit(fn, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(iters):
            flush.zero_()
            comm.Barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return max(comm.allgather(ts[len(ts) // 2]))

    rows = []
    shapes = [("qkv", H, 3 * H), ("fc1", H, 4 * H)]          # column-parallel: (in, out)
    rshapes = [("o", H, H), ("fc2", 4 * H, H)]               # row-parallel: (in, out)
    # ---- all-gather -> GEMM: forward of column layers, backward-dx of row layers ------------
    cases = [(f"ag_gemm fwd {nm}", din, dout // n, True) for nm, din, dout in shapes] + \
            [(f"ag_gemm bwd-dx {nm}", dout, din // n, False) for nm, din, dout in rshapes]
    only = os.environ.get("FB_ONLY", "")      # "ag" / "rs": one family only (kernel work in progress)
    if only == "rs":
        cases = []
    for name, K, Nl, b_kn in cases:
        op = F.AllGatherGemm(comm, M, K)
        x = (torch.randn(M, K // n, device="cuda") * 0.5).bfloat16()
        w = (torch.randn(K, Nl, device="cuda") * 0.05).bfloat16() if b_kn else (torch.randn(Nl, K, device="cuda") * 0.05).bfloat16()
        op.input_buffer().copy_(x)
        xin = op.input_buffer()
        out = torch.empty(M, Nl, device="cuda", dtype=torch.bfloat16)
        t_fused = timeit(lambda: op(xin, w, b_kn=b_kn, out=out))
        xfull = torch.randn(M, K, device="cuda").bfloat16()
        t_gemm = timeit(lambda: gemm(xfull, w, b_kn=b_kn, out=out))
        wt = w if b_kn else w.t()
        t_cublas = timeit(lambda: torch.matmul(xfull, wt, out=out))
        buf = torch.empty(n, M, K // n, device="cuda", dtype=torch.bfloat16)

        def nccl_ag():
            dist.all_gather_into_tensor(buf, x)
            return buf.permute(1, 0, 2).reshape(M, K)
        t_coll = timeit(nccl_ag)
        t_base = timeit(lambda: torch.matmul(nccl_ag(), wt, out=out))
        flops = 2.0 * M * K * Nl
        link = M * (K // n) * 2 * (n - 1)     # bytes this rank must receive
        rows.append(dict(op=name, M=M, N=Nl, K=K, fused_ms=t_fused, our_gemm_ms=t_gemm, cublas_ms=t_cublas,
                         nccl_coll_ms=t_coll, nccl_cublas_ms=t_base, flops=flops, link_bytes=link))
        if KT.enabled():        # CCMPI_KERNEL_TRACE=1: where the fused kernel's warp roles waited, next to the plain GEMM's
            flush.zero_(); comm.Barrier(); op(xin, w, b_kn=b_kn, out=out)
            rows[-1]["trace_fused"] = KT.summary(KT.last(op))
            tb = KT.new_buffer(); flush.zero_(); gemm(xfull, w, b_kn=b_kn, out=out, trace=tb)
            rows[-1]["trace_gemm"] = KT.summary(tb)
    # ---- GEMM -> reduce-scatter: forward of row layers, backward-dx of column layers --------
    cases = [(f"gemm_rs fwd {nm}", din // n, dout, True) for nm, din, dout in rshapes] + \
            [(f"gemm_rs bwd-dx {nm}", dout // n, din, False) for nm, din, dout in shapes]
    if only == "ag":
        cases = []
    for name, K, Nf, b_kn in cases:
        op = F.GemmReduceScatter(comm, M, Nf)
        a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
        w = (torch.randn(K, Nf, device="cuda") * 0.05).bfloat16() if b_kn else (torch.randn(Nf, K, device="cuda") * 0.05).bfloat16()
        out = torch.empty(M, Nf // n, device="cuda", dtype=torch.bfloat16)
        t_fused = timeit(lambda: op(a, w, b_kn=b_kn, out=out))
        full = torch.empty(M, Nf, device="cuda", dtype=torch.bfloat16)
        t_gemm = timeit(lambda: gemm(a, w, b_kn=b_kn, out=full))
        wt = w if b_kn else w.t()
        t_cublas = timeit(lambda: torch.matmul(a, wt, out=full))

        def nccl_rs(y):
            blocks = y.view(M, n, Nf // n).permute(1, 0, 2).contiguous()
            dist.reduce_scatter_tensor(out, blocks)
        t_coll = timeit(lambda: nccl_rs(full))
        t_base = timeit(lambda: nccl_rs(torch.matmul(a, wt, out=full)))
        flops = 2.0 * M * K * Nf
        link = M * (Nf // n) * 2 * (n - 1)
        rows.append(dict(op=name, M=M, N=Nf, K=K, fused_ms=t_fused, our_gemm_ms=t_gemm, cublas_ms=t_cublas,
                         nccl_coll_ms=t_coll, nccl_cublas_ms=t_base, flops=flops, link_bytes=link))
        if KT.enabled():
            flush.zero_(); comm.Barrier(); op(a, w, b_kn=b_kn, out=out)
            rows[-1]["trace_fused"] = KT.summary(KT.last(op))
            tb = KT.new_buffer(); flush.zero_(); gemm(a, w, b_kn=b_kn, out=full, trace=tb)
            rows[-1]["trace_gemm"] = KT.summary(tb)
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/fused_bench_{n}gpu.jsonl", "w") as f:
            for r in rows:
                ideal = max(r["flops"] / (PEAK_TFLOPS * 1e12), r["link_bytes"] / (LINK_GBS * 1e9)) * 1e3
                r["roofline_ms"] = round(ideal, 4)
                r["roofline_frac"] = round(ideal / r["fused_ms"], 3)
                r["speedup_vs_nccl_cublas"] = round(r["nccl_cublas_ms"] / r["fused_ms"], 3)
                for k in ("fused_ms", "our_gemm_ms", "cublas_ms", "nccl_coll_ms", "nccl_cublas_ms"):
                    r[k] = round(r[k], 4)
                r["n_gpus"] = n
                f.write(json.dumps(r) + "\n")
                print(f'{r["op"]:22s} fused {r["fused_ms"]:.3f}  gemm {r["our_gemm_ms"]:.3f}  cublas {r["cublas_ms"]:.3f}  '
                      f'nccl {r["nccl_coll_ms"]:.3f}  nccl+cublas {r["nccl_cublas_ms"]:.3f}  x{r["speedup_vs_nccl_cublas"]:.2f}  '
                      f'roofline {r["roofline_frac"]:.2f}', flush=True)
    comm.Barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
