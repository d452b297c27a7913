"""Kernel-time breakdown of one training step (torch.profiler, rank 0). Not a benchmark: numbers
under a profiler are only used as shares. Usage: [torchrun ...] profile_step.py ours|nccl"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

impl = sys.argv[1] if len(sys.argv) > 1 else "ours"
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
os.environ.setdefault("CCMPI_HEAP_MB", "6144")
os.environ.setdefault("CCMPI_TRANSPORT", "device")
from mpi4py import MPI  # noqa: E402
from model.func_impl import get_info  # noqa: E402

H, TOK = 4096, 8192
mp = 2 if world >= 2 else 1
dp = world // mp
comm = MPI.COMM_WORLD
mp_idx, dp_idx, mp_comm, dp_comm, _, _ = get_info(comm=comm, rank=rank, mp_size=mp, dp_size=dp, fc_layer="fc_q", in_dim=H, out_dim=H)
if impl == "ours":
    from collective_communication_mpi_b200.models import TrainStep, TransformerFCBlock
    step = TrainStep(TransformerFCBlock(mp_comm, TOK, H, 4, seed=1), dp_comm, lr=1e-4)
else:
    from nccl_baseline import NcclTrainStep
    step = NcclTrainStep(rank, world, mp, dp, TOK, H, 4, lr=1e-4)
x = (torch.randn(TOK, H // mp, device="cuda") * 0.5).bfloat16()
t = (torch.randn(TOK, H // mp, device="cuda") * 0.5).bfloat16()
for _ in range(3):
    step(x, t)
torch.cuda.synchronize()
comm.Barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        step(x, t)
    torch.cuda.synchronize()
comm.Barrier()
if rank == 0:
    rows = [(e.key, e.device_time_total / 3.0, e.count // 3) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/profile_{impl}_{world}gpu.txt", "w") as f:
        f.write(f"# {impl} {world} GPU: device time per step {tot/1000:.3f} ms (sum over kernels, profiler)\n")
        for k, us, c in rows[:60]:
            f.write(f"{us:10.1f} us  x{c:3d}  {k[:110]}\n")
    print(open(f"gpurun_out/profile_{impl}_{world}gpu.txt").read())
if impl == "nccl":
    import torch.distributed as dist
    dist.destroy_process_group()
