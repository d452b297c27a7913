#!/bin/bash
# One ncu --set full capture of the top kernel (CTA-pair tcgen05 GEMM, fc1 forward shape) and one of
# the one-shot all-reduce with 2 emulated ranks. Each command first runs clean without ncu.
mkdir -p gpurun_out
cat > /tmp/one_gemm.py <<'PY'
import sys, torch
sys.path.insert(0, ".")
from collective_communication_mpi_b200.ops.gemm import gemm
M, N, K = 8192, 8192, 4096
a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16(); b = (torch.randn(K, N, device="cuda") * 0.5).bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(4): gemm(a, b, b_kn=True, out=out)
torch.cuda.synchronize(); print("ok", out.float().abs().mean().item())
PY
cat > /tmp/one_ar.py <<'PY'
import sys, torch
sys.path.insert(0, ".")
from collective_communication_mpi_b200 import _native as N
from collective_communication_mpi_b200.ops import collectives as CO
from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
heap = SymmetricHeap.create_emulated(2, 256 << 20); comms = [make_dev_comm(heap, [0, 1], r, 0) for r in range(2)]
cnt = 16 << 20
s, d, t = heap.alloc(cnt * 2), heap.alloc(cnt * 2), heap.alloc(cnt * 2)
for _ in range(4): CO.allreduce(comms, N.BF16, N.SUM, s, d, cnt, tmp_off=t, algo=N.AR_TWOSHOT)
torch.cuda.synchronize(); print("ok")
PY
python /tmp/one_gemm.py > gpurun_out/plain_gemm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel_2cta -s 2 -c 1 -f -o gpurun_out/prof_gemm_2cta python /tmp/one_gemm.py > gpurun_out/ncu_gemm.log 2>&1
tail -2 gpurun_out/ncu_gemm.log
python /tmp/one_ar.py > gpurun_out/plain_ar.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:allreduce_twoshot -s 2 -c 1 -f -o gpurun_out/prof_ar_twoshot python /tmp/one_ar.py > gpurun_out/ncu_ar.log 2>&1
tail -2 gpurun_out/ncu_ar.log
ls -la gpurun_out/*.ncu-rep
