#!/bin/bash
# Multi-GPU bring-up: runtime bootstrap (heap exchange, multicast), device-plane tests, sweep.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
{
echo "== $NG GPUs =="; nvidia-smi topo -m | head -12
python - <<'PY'
import torch
n = torch.cuda.device_count()
print("p2p:", [[int(torch.cuda.can_device_access_peer(i, j)) if i != j else 1 for j in range(n)] for i in range(n)])
PY
echo "== bootstrap =="
timeout 120 python -m collective_communication_mpi_b200.launch -n $NG --transport device --tag-output python - <<'PY'
from mpi4py import MPI
import torch
c = MPI.COMM_WORLD
r = c.Get_rank(); torch.cuda.set_device(r)
d = c.device_comm()
print("rank", r, "heap MB", d.heap.bytes >> 20, "mc", d.has_mc, "mc_supported", c._world.mc_supported, "err", c._world.device_error)
x = c.symm_empty(1 << 20, torch.float32); x.fill_(r + 1); y = c.symm_empty(1 << 20, torch.float32)
from collective_communication_mpi_b200.ops import collectives as CO
from collective_communication_mpi_b200 import _native as N
h = d.heap
for name, algo in (("oneshot", N.AR_ONESHOT), ("twoshot", N.AR_TWOSHOT)) + ((("nvls", N.AR_NVLS),) if d.has_mc else ()):
    y.zero_(); torch.cuda.synchronize(); c.Barrier()
    t = c.symm_empty(1 << 20, torch.float32)
    CO.allreduce(d.comms, N.F32, N.SUM, h.offset_of(x), h.offset_of(y), 1 << 20, tmp_off=h.offset_of(t), algo=algo)
    torch.cuda.synchronize()
    n = c.Get_size()
    print("rank", r, name, "ok" if bool((y == n * (n + 1) / 2).all()) else f"WRONG {y[:4].tolist()}")
c.Barrier()
PY
echo "== pytest multirank =="
timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q 2>&1 | tail -15
} > gpurun_out/check2_${NG}gpu.txt 2>&1
tail -40 gpurun_out/check2_${NG}gpu.txt
echo "== sweep =="
timeout 600 python -m collective_communication_mpi_b200.launch -n $NG --transport device python benchmarks/sweep.py --max-mb ${SWEEP_MB:-256} 2>&1 | tail -100 > gpurun_out/sweep_${NG}gpu.log
tail -5 gpurun_out/sweep_${NG}gpu.log
