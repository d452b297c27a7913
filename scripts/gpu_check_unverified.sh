#!/bin/bash
# First GPU call of the next round: everything that was written after round 1's GPU budget was
# spent, each file under its own timeout so one hang cannot take the rest (or the box) with it.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_check_unverified.sh'            (1 GPU)
#   gpurun --gpus 2 --timeout 1500 -- 'bash scripts/gpu_check_unverified.sh'   (adds the multi-process checks)
# Results: gpurun_out/unverified_<N>gpu.txt (summary) and one log per test file.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
OUT=gpurun_out/unverified_${NG}gpu.txt
export CCMPI_TEST_UNVERIFIED=1
: > "$OUT"
run() {   # run <label> <timeout> <command...>
    local label=$1 limit=$2; shift 2
    local log=gpurun_out/unverified_${label}_${NG}gpu.log
    timeout "$limit" "$@" > "$log" 2>&1
    local rc=$?
    echo "[$label] rc=$rc  $(tail -1 "$log")" | tee -a "$OUT"
}
run regular_suite 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_unverified_fp8.py \
    --deselect tests/test_gpu_unverified_seq.py --deselect tests/test_gpu_unverified_elementwise.py --deselect tests/test_gpu_unverified_p2p.py --deselect tests/test_gpu_unverified_sgd.py --deselect tests/test_gpu_unverified_graphs.py --deselect tests/test_gpu_unverified_tight.py
run utils        120 python -m pytest tests/test_gpu_utils.py -x -q
run elementwise  300 python -m pytest tests/test_gpu_unverified_elementwise.py -q
run tight        300 python -m pytest tests/test_gpu_unverified_tight.py -q
run graphs       120 python -m pytest tests/test_gpu_unverified_graphs.py -q
run sgd          300 python -m pytest tests/test_gpu_unverified_sgd.py -q
run fp8          300 python -m pytest tests/test_gpu_unverified_fp8.py -q
run seq_emulated 600 python -m pytest tests/test_gpu_unverified_seq.py -q -k "not across_processes"
if [ "$NG" -ge 2 ]; then
    run seq_multiproc 300 python -m pytest tests/test_gpu_unverified_seq.py -q -k across_processes
    run p2p_device    300 python -m pytest tests/test_gpu_unverified_p2p.py -q
    run put_signal    300 python -m collective_communication_mpi_b200.launch -n "$NG" --transport device python tests/workers/w_collectives.py --cuda
fi
run extras_bench 600 python benchmarks/extras_bench.py
cp gpurun_out/unverified_extras_bench_${NG}gpu.log gpurun_out/extras_bench_${NG}gpu.jsonl 2>/dev/null
# the switch that would become the default once the elementwise kernels are verified
CCMPI_FUSED_ELEMENTWISE=0 run bench_eager 300 python bench.py --steps 10 --warmup 3
CCMPI_FUSED_ELEMENTWISE=1 run bench_fused 300 python bench.py --steps 10 --warmup 3
CCMPI_FUSED_ELEMENTWISE=1 CCMPI_FUSED_SGD=1 run bench_fused_sgd 300 python bench.py --steps 10 --warmup 3
if [ "$NG" -ge 2 ]; then
    run sweep_tight 600 python -m collective_communication_mpi_b200.launch -n "$NG" --transport device python benchmarks/sweep.py --tight --no-nccl --min-kb 1024 --max-mb 256
    run seq_bench 600 python -m collective_communication_mpi_b200.launch -n "$NG" --transport device python benchmarks/seq_bench.py
    # end-to-end (H2D-bound at N >= 4 in round 1): does running each rank next to its GPU help?
    P=29600
    for bind in 0 1; do
        P=$((P + 1))
        CCMPI_NUMA_BIND=$bind run bench_e2e_numa$bind 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NG" \
            --master-addr 127.0.0.1 --master-port $P bench.py --gpus "$NG" --steps 10 --warmup 3
    done
fi
cat "$OUT"
# how long one reference-arm CPU step takes on this box (sets the default wall budget of the arm)
CCMPI_REF_BUDGET_S=${REF_PROBE_BUDGET_S:-70} run reference_arm 400 python bench.py --impl reference --gpus 1 --steps 2 --warmup 3
nproc >> "$OUT"; cat "$OUT"
