import os, sys, time, torch
sys.path.insert(0, "."); sys.path.insert(0, "benchmarks")
os.environ.setdefault("CCMPI_TRANSPORT", "device"); os.environ.setdefault("CCMPI_HEAP_MB", "4096")
from mpi4py import MPI
from model.func_impl import get_info
from collective_communication_mpi_b200.models import TrainStep, TransformerFCBlock
from collective_communication_mpi_b200.utils.data import AsyncScalarReader, DevicePrefetcher
from nccl_baseline import NcclTrainStep
comm = MPI.COMM_WORLD
_, _, mpc, dpc, _, _ = get_info(comm=comm, rank=0, mp_size=1, dp_size=1, fc_layer="fc_q", in_dim=4096, out_dim=4096)
TOK, H = 8192, 4096
for impl in ("ours", "nccl"):
    step = TrainStep(TransformerFCBlock(mpc, TOK, H, 4, seed=1), dpc, lr=1e-4) if impl == "ours" else NcclTrainStep(0, 1, 1, 1, TOK, H, 4)
    hx = [(torch.randn(TOK, H) * 0.5).bfloat16().pin_memory() for _ in range(4)]
    ht = [(torch.randn(TOK, H) * 0.5).bfloat16().pin_memory() for _ in range(4)]
    dx = [h.cuda() for h in hx]; dt = [h.cuda() for h in ht]
    def run(mode, steps=20):
        pf, rd = DevicePrefetcher([hx[0], ht[0]]), AsyncScalarReader()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record()
        if "h2d" in mode: pf.submit((hx[0], ht[0]))
        cpu = 0.0
        for i in range(steps):
            if "h2d" in mode:
                xb, tb = pf.get()
                if i + 1 < steps: pf.submit((hx[(i + 1) % 4], ht[(i + 1) % 4]))
            else:
                xb, tb = dx[i % 4], dt[i % 4]
            c0 = time.perf_counter(); loss = step(xb, tb); cpu += time.perf_counter() - c0
            if "h2d" in mode: pf.release()
            if "loss" in mode:
                rd.push(loss)
                if rd.pending > 2: rd.pop()
        while rd.pending: rd.pop()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps, cpu / steps * 1e3
    for _ in range(5): step(dx[0], dt[0])
    for mode in ("device", "loss", "h2d", "h2d+loss"):
        run(mode, 4)
        ms, cpu = run(mode)
        print(f"{impl:5s} {mode:9s} {ms:.3f} ms/step   cpu launch {cpu:.3f} ms/step", flush=True)
