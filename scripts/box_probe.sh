#!/bin/bash
# One-shot description of the GPU box. Output goes to gpurun_out/box_probe.txt
mkdir -p gpurun_out
{
  echo "== nvidia-smi =="; nvidia-smi --query-gpu=index,name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv
  echo "== topo =="; nvidia-smi topo -m
  echo "== nvlink =="; nvidia-smi nvlink -s 2>&1 | head -30
  echo "== cpu =="; nproc; lscpu | grep -E "Model name|Socket|NUMA" ; free -g | head -2
  echo "== shm =="; df -h /dev/shm
  echo "== mpi? =="; which mpirun mpiexec 2>&1; python -c "import mpi4py" 2>&1 | tail -1
  echo "== repo layout =="; ls -la | head -40
  echo "== probe =="; python - <<'PY'
import ctypes, sys
sys.path.insert(0, '.')
from collective_communication_mpi_b200 import _native as N
lib = N.device_lib()
out = (ctypes.c_int * 8)()
import torch
n = torch.cuda.device_count()
for d in range(n):
    lib.ccmpi_probe(d, out)
    print("device", d, dict(zip(["ndev", "cc_major", "cc_minor", "sms", "multicast", "vmm", "posix_fd", "gran"], list(out))))
if n > 1:
    print("p2p matrix:", [[int(torch.cuda.can_device_access_peer(i, j)) if i != j else 1 for j in range(n)] for i in range(n)])
PY
} > gpurun_out/box_probe.txt 2>&1
tail -40 gpurun_out/box_probe.txt
