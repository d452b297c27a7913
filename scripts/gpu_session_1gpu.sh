#!/bin/bash
# Routine 1-GPU session: the GPU test-suite, a kernel-time breakdown of the training step (torch.profiler,
# shares only), and both GPU arms of the headline bench. Results under gpurun_out/.
mkdir -p gpurun_out
export CCMPI_DEVICE_TIMEOUT_S=${CCMPI_DEVICE_TIMEOUT_S:-20}
OUT=gpurun_out/session_1gpu.txt
: > "$OUT"
run() { local label=$1 limit=$2; shift 2
    timeout "$limit" "$@" > gpurun_out/s1_${label}.log 2>&1
    echo "[$label] rc=$?  $(tail -1 gpurun_out/s1_${label}.log)" | tee -a "$OUT"; }
run tests 900 python -m pytest tests -x -q -m gpu
run profile_ours 300 python benchmarks/profile_step.py ours
run bench_ours 300 python bench.py --steps 20 --warmup 5
run bench_nccl 300 python bench.py --impl nccl --steps 20 --warmup 5
for extra in "$@"; do run extra_$(echo "$extra" | tr -c 'a-zA-Z0-9' _ | cut -c1-40) 600 bash -c "$extra"; done
cat "$OUT"
