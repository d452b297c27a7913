"""Token-axis (sequence-parallel) fused ops vs NCCL + cuBLAS, per FC shape.

On the token axis NCCL needs no layout copies (``all_gather_into_tensor`` / ``reduce_scatter_tensor``
work on contiguous row blocks), so this is the baseline at its best. Device-timed (CUDA events),
median of 10, max over ranks, L2 flushed between iterations.

    python -m collective_communication_mpi_b200.launch -n 2 --transport device python benchmarks/seq_bench.py

The kernels measured here have not run on hardware yet (see ops/fused_seq.py); results are checked
against the baseline's output before anything is timed.
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
from collective_communication_mpi_b200.ops.fused_seq import AllGatherGemmM, GemmReduceScatterM  # noqa: E402


def main():
    comm = MPI.COMM_WORLD
    rank, n = comm.Get_rank(), comm.Get_size()
    torch.cuda.set_device(rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29577")
    dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    H = int(os.environ.get("FB_HIDDEN", "4096"))
    M = int(os.environ.get("FB_TOKENS", "8192"))
    rows = M // n
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

    def rnd(*shape, seed):
        g = torch.Generator(device="cuda").manual_seed(seed)
        return (torch.randn(*shape, device="cuda", generator=g) * 0.1).bfloat16()

    out = []
    # all-gather (rows) -> GEMM: column-parallel forward in the sequence-parallel layout
    for name, K, N in (("qkv", H, 3 * H // n), ("fc1", H, 4 * H // n)):
        op = AllGatherGemmM(comm, M, K)
        x, w = rnd(rows, K, seed=rank), rnd(K, N, seed=100 + rank)
        op.input_buffer().copy_(x)
        xs = op.input_buffer()
        full = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)

        def base():
            dist.all_gather_into_tensor(full, x)
            return full @ w

        y, ref = op(xs, w), base()
        torch.cuda.synchronize()
        err = (y.float() - ref.float()).abs().max().item()
        assert err < 0.05 * max(1.0, ref.float().abs().max().item()), f"ag_gemm_m {name}: mismatch {err}"
        t_f, t_b = timeit(lambda: op(xs, w)), timeit(base)
        out.append(dict(op=f"ag_gemm_m {name}", M=M, N=N, K=K, fused_ms=round(t_f, 4), nccl_cublas_ms=round(t_b, 4),
                        speedup=round(t_b / t_f, 3)))
    # GEMM -> reduce-scatter (rows): row-parallel forward in the sequence-parallel layout
    for name, K, N in (("o", H // n, H), ("fc2", 4 * H // n, H)):
        op = GemmReduceScatterM(comm, M, N)
        a, w = rnd(M, K, seed=200 + rank), rnd(K, N, seed=300 + rank)
        red = torch.empty(rows, N, device="cuda", dtype=torch.bfloat16)

        def base():
            dist.reduce_scatter_tensor(red, a @ w)
            return red

        y, ref = op(a, w, b_kn=True), base()
        torch.cuda.synchronize()
        err = (y.float() - ref.float()).abs().max().item()
        assert err < 0.05 * max(1.0, ref.float().abs().max().item()), f"gemm_rs_m {name}: mismatch {err}"
        t_f, t_b = timeit(lambda: op(a, w, b_kn=True)), timeit(base)
        out.append(dict(op=f"gemm_rs_m {name}", M=M, N=N, K=K, fused_ms=round(t_f, 4), nccl_cublas_ms=round(t_b, 4),
                        speedup=round(t_b / t_f, 3)))
    if rank == 0:
        for r in out:
            print(json.dumps(dict(r, n_gpus=n, hidden=H)), flush=True)
    comm.Barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
