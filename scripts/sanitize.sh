#!/bin/bash
# Race / memory checking of the single-GPU kernels ("race detection / sanitizers").
# compute-sanitizer sees one GPU only, so it is pointed at (a) the plain tcgen05 GEMM and (b) the
# collectives and fused kernels with EMULATED ranks (all ranks' CTAs inside one cooperative launch).
# The profiling guide's rule applies: ONE sanitizer tool per gpurun call; pick it with $1.
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck'
# Peer-to-peer flag protocols across real GPUs are out of any sanitizer's sight; those are covered by
# the repeated-launch / changing-payload tests in tests/test_gpu_multirank.py and the device watchdog.
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
cat > /tmp/san_case.py <<'PY'
import sys, torch
sys.path.insert(0, ".")
from collective_communication_mpi_b200 import _native as N
from collective_communication_mpi_b200.ops import collectives as CO
from collective_communication_mpi_b200.ops.gemm import gemm
from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
a = (torch.randn(384, 320, device="cuda") * 0.5).bfloat16(); b = (torch.randn(256, 320, device="cuda") * 0.5).bfloat16()
out = gemm(a, b, out_dtype=torch.float32)
heap = SymmetricHeap.create_emulated(2, 8 << 20); comms = [make_dev_comm(heap, [0, 1], r, 0) for r in range(2)]
s, d, t = heap.alloc(1 << 16), heap.alloc(1 << 16), heap.alloc(1 << 16)
for algo in (N.AR_ONESHOT, N.AR_TWOSHOT):
    CO.allreduce(comms, N.F32, N.SUM, s, d, 1 << 12, tmp_off=t, algo=algo)
CO.allgather(comms, s, d, 1, 1 << 12, 0, 0); CO.alltoall(comms, s, d, 1 << 12); CO.barrier(comms)
torch.cuda.synchronize(); print("sanitizer case ok", float(out.abs().mean()))
PY
python /tmp/san_case.py > gpurun_out/sanitize_plain.log 2>&1 && \
compute-sanitizer --tool "$TOOL" --error-exitcode 7 python /tmp/san_case.py > gpurun_out/sanitize_${TOOL}.log 2>&1
echo "exit $?"; tail -5 gpurun_out/sanitize_${TOOL}.log
