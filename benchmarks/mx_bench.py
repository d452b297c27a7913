"""Block-scaled fp8 (MXFP8) fused ops vs their baselines, forward products of the FC layers (BASELINE config 3).

Per shape: our fused all-gather->GEMM / GEMM->reduce-scatter in MXFP8 (quantisation of the activation included
in the timed call) | the same op in bf16 (fused) | NCCL collective + torch._scaled_mm (rowwise-scaled fp8, the
library's closest block-scaled-like mode that exists for arbitrary shapes; falls back to per-tensor scales) .
Device-timed, max over ranks, L2 flushed between iterations.

    python -m collective_communication_mpi_b200.launch -n 8 python benchmarks/mx_bench.py
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
from collective_communication_mpi_b200.ops import gemm_fp8 as Q  # noqa: E402


def main():
    comm = MPI.COMM_WORLD
    rank, n = comm.Get_rank(), comm.Get_size()
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    H = int(os.environ.get("FB_HIDDEN", "8192" if n >= 8 else "4096"))
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

    def scaled_mm(a_bf16, wq, wscale):
        """library fp8 GEMM: per-tensor-scaled e4m3 through torch._scaled_mm (cuBLASLt)."""
        amax = a_bf16.abs().amax().float().clamp_min(1e-12)
        sa = amax / 448.0
        aq = (a_bf16.float() / sa).to(torch.float8_e4m3fn)
        return torch._scaled_mm(aq, wq.t(), scale_a=sa.reshape(1), scale_b=wscale.reshape(1), out_dtype=torch.bfloat16)

    rows = []
    # ---- all-gather -> GEMM (column-parallel forward): qkv and fc1 ----
    for name, K, Nl in (("ag_gemm fwd qkv", H, 3 * H // n), ("ag_gemm fwd fc1", H, 4 * H // n)):
        x = (torch.randn(M, K // n, device="cuda") * 0.5).bfloat16()
        w = (torch.randn(K, Nl, device="cuda") * 0.05).bfloat16()
        mx, bf = Q.AllGatherGemmMX(comm, M, K), F.AllGatherGemm(comm, M, K)
        wq = Q.QuantizedWeight(w)
        out = torch.empty(M, Nl, device="cuda", dtype=torch.bfloat16)
        t_mx = timeit(lambda: mx(x, wq, out=out))
        y_mx = out.clone()
        t_bf = timeit(lambda: bf(x, w, b_kn=True, out=out))
        err = float((y_mx.float() - out.float()).abs().max() / out.float().abs().max())
        wamax = w.abs().amax().float()
        w8 = (w.t().float() / (wamax / 448.0)).to(torch.float8_e4m3fn).contiguous()       # [N, K]
        buf = torch.empty(n, M, K // n, device="cuda", dtype=torch.bfloat16)

        def lib():
            dist.all_gather_into_tensor(buf, x)
            return scaled_mm(buf.permute(1, 0, 2).reshape(M, K), w8, wamax / 448.0)
        t_lib = timeit(lib)
        rows.append(dict(op=name, M=M, N=Nl, K=K, mxfp8_fused_ms=t_mx, bf16_fused_ms=t_bf, nccl_scaled_mm_ms=t_lib,
                         rel_diff_vs_bf16=err, tflops_mx=2.0 * M * K * Nl / t_mx / 1e9))
    # ---- GEMM -> reduce-scatter (row-parallel forward): o and fc2 ----
    for name, K, Nf in (("gemm_rs fwd o", H // n, H), ("gemm_rs fwd fc2", 4 * H // n, H)):
        a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
        w = (torch.randn(K, Nf, device="cuda") * 0.05).bfloat16()
        mx, bf = Q.GemmReduceScatterMX(comm, M, Nf), F.GemmReduceScatter(comm, M, Nf)
        wq = Q.QuantizedWeight(w)
        out = torch.empty(M, Nf // n, device="cuda", dtype=torch.bfloat16)
        t_mx = timeit(lambda: mx(a, wq, out=out))
        y_mx = out.clone()
        t_bf = timeit(lambda: bf(a, w, b_kn=True, out=out))
        err = float((y_mx.float() - out.float()).abs().max() / out.float().abs().max())
        wamax = w.abs().amax().float()
        w8 = (w.t().float() / (wamax / 448.0)).to(torch.float8_e4m3fn).contiguous()

        def lib():
            y = scaled_mm(a, w8, wamax / 448.0)
            blocks = y.view(M, n, Nf // n).permute(1, 0, 2).contiguous()
            dist.reduce_scatter_tensor(out, blocks)
        t_lib = timeit(lib)
        rows.append(dict(op=name, M=M, N=Nf, K=K, mxfp8_fused_ms=t_mx, bf16_fused_ms=t_bf, nccl_scaled_mm_ms=t_lib,
                         rel_diff_vs_bf16=err, tflops_mx=2.0 * M * K * Nf / t_mx / 1e9))
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/mx_bench_{n}gpu.jsonl", "w") as f:
            for r in rows:
                r = {k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}
                r["n_gpus"], r["hidden"] = n, H
                r["speedup_vs_nccl_scaled_mm"] = round(r["nccl_scaled_mm_ms"] / r["mxfp8_fused_ms"], 3)
                r["speedup_vs_bf16_fused"] = round(r["bf16_fused_ms"] / r["mxfp8_fused_ms"], 3)
                f.write(json.dumps(r) + "\n")
                print(json.dumps(r), flush=True)
    comm.Barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
