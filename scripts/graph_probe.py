"""Is the 1-GPU training step launch-bound anywhere? Time the step launched from Python, then the same
step captured once into a CUDA graph and replayed (no Python, no per-kernel launch cost), and the sum of
its kernels' own times from the profiler. If replay ~= eager, the gaps between kernels are not launch
overhead. Prints three ms/step numbers."""
import os
import sys

sys.path.insert(0, ".")
os.environ.setdefault("CCMPI_TRANSPORT", "device")
os.environ.setdefault("CCMPI_HEAP_MB", "4096")
import torch
from mpi4py import MPI
from model.func_impl import get_info
from collective_communication_mpi_b200.models import TrainStep, TransformerFCBlock

comm = MPI.COMM_WORLD
_, _, mpc, dpc, _, _ = get_info(comm=comm, rank=0, mp_size=1, dp_size=1, fc_layer="fc_q", in_dim=4096, out_dim=4096)
TOK, H = 8192, 4096
step = TrainStep(TransformerFCBlock(mpc, TOK, H, 4, seed=1), dpc, lr=1e-4)
x = (torch.randn(TOK, H, device="cuda") * 0.5).bfloat16()
t = (torch.randn(TOK, H, device="cuda") * 0.5).bfloat16()


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


eager = timed(lambda: step(x, t))
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(3):
        step(x, t)
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    loss = step(x, t)
replay = timed(g.replay)
print(f"eager {eager:.3f} ms/step   graph replay {replay:.3f} ms/step   loss {float(loss):.4f}", flush=True)
