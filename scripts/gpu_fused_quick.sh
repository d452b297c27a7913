#!/bin/bash
# Quick check of the fused GEMM+collective kernels while they are being worked on: the emulated-rank
# numerics tests (1 GPU) and the fused-op benchmark with the device trace on the box's GPUs.
#   gpurun --gpus 2 -- 'FB_ONLY=ag bash scripts/gpu_fused_quick.sh'
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
export CCMPI_DEVICE_TIMEOUT_S=${CCMPI_DEVICE_TIMEOUT_S:-20}
timeout 600 python -m pytest tests/test_gpu_fused_emulated.py tests/test_gpu_epilogue.py tests/test_gpu_seq.py -x -q -k "not across_processes and not gathered_output" > gpurun_out/q_tests.log 2>&1
echo "[tests] rc=$? $(tail -1 gpurun_out/q_tests.log)"
CCMPI_KERNEL_TRACE=1 timeout 600 python -m collective_communication_mpi_b200.launch -n "$NG" --transport device python benchmarks/fused_bench.py > gpurun_out/q_fused.log 2>&1
echo "[fused] rc=$?"; grep -v "^\[" gpurun_out/q_fused.log | tail -9
cp gpurun_out/fused_bench_${NG}gpu.jsonl gpurun_out/fused_bench_${NG}gpu_trace.jsonl 2>/dev/null
