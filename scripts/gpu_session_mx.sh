#!/bin/bash
# Block-scaled fp8 bring-up + the data-parallel optimizer paths on a small box.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
export CCMPI_DEVICE_TIMEOUT_S=${CCMPI_DEVICE_TIMEOUT_S:-20}
run() { local label=$1 limit=$2; shift 2
    timeout "$limit" "$@" > gpurun_out/mx_${label}.log 2>&1
    echo "[$label] rc=$?  $(tail -1 gpurun_out/mx_${label}.log | cut -c1-600)"; }
run tests_mx 600 python -m pytest tests/test_gpu_mx_fused.py tests/test_gpu_fp8.py -q --maxfail=5
run tests_opt 600 python -m pytest tests/test_gpu_sgd.py tests/test_gpu_epilogue.py tests/test_gpu_fused_emulated.py -q --maxfail=5
if [ "$NG" -ge 2 ]; then
    run dp_update 300 python -m pytest tests/test_gpu_multirank.py -q -k "dp_allreduce_fused or match_dense_model" --maxfail=3
    run bench_dp2 400 env CCMPI_BENCH_MP=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29911 bench.py --gpus 2 --steps 10 --warmup 3
    run bench_mp2 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29912 bench.py --gpus 2 --steps 20 --warmup 5
    run mx_bench 400 python -m collective_communication_mpi_b200.launch -n "$NG" --transport device python benchmarks/mx_bench.py
fi
