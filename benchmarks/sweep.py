"""Collective sweep: all-reduce (one-shot / two-shot / NVLS) /
all-gather / reduce-scatter / all-to-all from 1 KB to 1 GB, device-timed, max over ranks,
against NCCL through torch.distributed on the same buffers. Bus bandwidth uses the nccl-tests
convention: all-reduce 2(n-1)/n * S/t, the others (n-1)/n * S/t, S = per-rank payload
(all-reduce) or total gathered/scattered bytes.

    python -m collective_communication_mpi_b200.launch -n 8 python benchmarks/sweep.py --max-mb 1024

Writes one JSON line per (op, impl, size) to gpurun_out/sweep_<n>gpu.jsonl (rank 0).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from mpi4py import MPI  # noqa: E402
from collective_communication_mpi_b200 import _native as N  # noqa: E402
from collective_communication_mpi_b200.ops import collectives as CO  # noqa: E402


def timeit(fn, iters, warmup, comm):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    comm.Barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return max(comm.allgather(ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-kb", type=int, default=1)
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--blocks", type=int, default=0, help="max blocks for our kernels (0 = default)")
    ap.add_argument("--out", default=None)
    ap.add_argument("--tight", action="store_true",
                    help="also time the group-size-specialised kernels of csrc/collectives_tight.cu by name")
    args = ap.parse_args()
    os.environ.setdefault("CCMPI_HEAP_MB", str(max(2048, 3 * args.max_mb + 512)))
    comm = MPI.COMM_WORLD
    rank, n = comm.Get_rank(), comm.Get_size()
    torch.cuda.set_device(rank)
    dt = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    cc = N.BF16 if args.dtype == "bf16" else N.F32
    es = 2 if args.dtype == "bf16" else 4
    dev = comm.device_comm()
    heap = dev.heap
    max_bytes = args.max_mb << 20
    src = comm.symm_empty(max_bytes // es, dt)
    dst = comm.symm_empty(max_bytes // es, dt)
    tmp = comm.symm_empty(max_bytes // es, dt)
    src.normal_()
    so, do, to = heap.offset_of(src), heap.offset_of(dst), heap.offset_of(tmp)
    use_nccl = not args.no_nccl
    if use_nccl:
        import torch.distributed as dist
        dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    out_path = args.out or os.path.join("gpurun_out", f"sweep_{n}gpu.jsonl")
    rows = []
    sizes = []
    s = args.min_kb << 10
    while s <= max_bytes:
        sizes.append(s)
        s *= 4
    if sizes[-1] != max_bytes:
        sizes.append(max_bytes)
    mb = args.blocks

    def rec(op, impl, S, ms, factor):
        bus = factor * S / (ms * 1e-3) / 1e9
        rows.append({"op": op, "impl": impl, "bytes": S, "ms": round(ms, 5), "busbw_GBs": round(bus, 2), "n": n,
                     "dtype": args.dtype})
        if rank == 0:
            print(json.dumps(rows[-1]), flush=True)

    for S in sizes:
        cnt = S // es
        iters = 200 if S <= (1 << 20) else (50 if S <= (64 << 20) else 10)
        warm = 5
        f_ar = 2 * (n - 1) / n
        f_x = (n - 1) / n
        algos = [("oneshot", N.AR_ONESHOT)] if S <= (8 << 20) else []
        algos.append(("twoshot", N.AR_TWOSHOT))
        if dev.has_mc:
            algos.append(("nvls", N.AR_NVLS))
        for name, algo in algos:
            ms = timeit(lambda: CO.allreduce(dev.comms, cc, N.SUM, so, do, cnt, tmp_off=to, algo=algo, max_blocks=mb, tight=False), iters, warm, comm)
            rec("allreduce", name, S, ms, f_ar)
        ms = timeit(lambda: CO.allreduce(dev.comms, cc, N.SUM, so, do, cnt, tmp_off=to, algo=N.AR_AUTO, max_blocks=mb), iters, warm, comm)
        rec("allreduce", "auto", S, ms, f_ar)      # what a caller who names no algorithm gets
        if args.tight:
            ms = timeit(lambda: CO.allreduce(dev.comms, cc, N.SUM, so, do, cnt, tmp_off=to, algo=N.AR_TWOSHOT, max_blocks=mb, tight=True), iters, warm, comm)
            rec("allreduce", "twoshot_tight", S, ms, f_ar)
        per = S // n // 16 * 16
        if per >= 16:
            ms = timeit(lambda: CO.allgather(dev.comms, so, do, 1, per, 0, 0, use_mc=False, max_blocks=mb), iters, warm, comm)
            rec("allgather", "push_p2p", per * n, ms, f_x)
            if dev.has_mc:
                ms = timeit(lambda: CO.allgather(dev.comms, so, do, 1, per, 0, 0, use_mc=2, max_blocks=mb), iters, warm, comm)
                rec("allgather", "push_mc", per * n, ms, f_x)
            ms = timeit(lambda: CO.reduce_scatter(dev.comms, cc, N.SUM, so, do, 1, per // es, 0, 0, use_mc=False, max_blocks=mb, tight=False), iters, warm, comm)
            rec("reduce_scatter", "pull_p2p", per * n, ms, f_x)
            if args.tight:
                ms = timeit(lambda: CO.reduce_scatter(dev.comms, cc, N.SUM, so, do, 1, per // es, 0, 0, use_mc=False, max_blocks=mb, tight=True), iters, warm, comm)
                rec("reduce_scatter", "pull_p2p_tight", per * n, ms, f_x)
            if dev.has_mc:
                ms = timeit(lambda: CO.reduce_scatter(dev.comms, cc, N.SUM, so, do, 1, per // es, 0, 0, use_mc=3, max_blocks=mb), iters, warm, comm)
                rec("reduce_scatter", "nvls", per * n, ms, f_x)
            ms = timeit(lambda: CO.allgather(dev.comms, so, do, 1, per, 0, 0, use_mc=True, max_blocks=mb), iters, warm, comm)
            rec("allgather", "auto", per * n, ms, f_x)
            ms = timeit(lambda: CO.reduce_scatter(dev.comms, cc, N.SUM, so, do, 1, per // es, 0, 0, use_mc=True, max_blocks=mb), iters, warm, comm)
            rec("reduce_scatter", "auto", per * n, ms, f_x)
            ms = timeit(lambda: CO.alltoall(dev.comms, so, do, per, max_blocks=mb), iters, warm, comm)
            rec("alltoall", "push", per * n, ms, f_x)
        if use_nccl:
            import torch.distributed as dist
            x = src[:cnt]
            y = dst[:cnt]
            ms = timeit(lambda: dist.all_reduce(y), iters, warm, comm)
            rec("allreduce", "nccl", S, ms, f_ar)
            if per >= 16:
                pe = per // es
                ms = timeit(lambda: dist.all_gather_into_tensor(dst[:pe * n], src[:pe]), iters, warm, comm)
                rec("allgather", "nccl", per * n, ms, f_x)
                ms = timeit(lambda: dist.reduce_scatter_tensor(dst[:pe], src[:pe * n]), iters, warm, comm)
                rec("reduce_scatter", "nccl", per * n, ms, f_x)
                ms = timeit(lambda: dist.all_to_all_single(dst[:pe * n], src[:pe * n]), iters, warm, comm)
                rec("alltoall", "nccl", per * n, ms, f_x)
    if rank == 0:
        os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
        with open(out_path, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    comm.Barrier()
    if use_nccl:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
