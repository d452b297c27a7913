#!/bin/bash
# Routine multi-GPU session (gpurun --gpus N, N = 2, 4 or 8): the GPU test-suite with its multi-process
# tests, the headline bench on N GPUs (both GPU arms), the collective sweep, the fused-op benchmark.
# Every step has its own timeout; results under gpurun_out/ with an _<N>gpu suffix.
#   SKIP="sweep fused"  leaves steps out;  EXTRA="cmd"  appends one more step.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
export CCMPI_DEVICE_TIMEOUT_S=${CCMPI_DEVICE_TIMEOUT_S:-20}

OUT=gpurun_out/session_${NG}gpu.txt
: > "$OUT"
P=29700
run() { local label=$1 limit=$2; shift 2
    case " $SKIP " in *" $label "*) echo "[$label] skipped" | tee -a "$OUT"; return;; esac
    timeout "$limit" "$@" > gpurun_out/s${NG}_${label}.log 2>&1
    echo "[$label] rc=$?  $(tail -1 gpurun_out/s${NG}_${label}.log | cut -c1-900)" | tee -a "$OUT"; }
torchrun_() { P=$((P + 1)); python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NG" --master-addr 127.0.0.1 --master-port $P "$@"; }
launch_() { python -m collective_communication_mpi_b200.launch -n "$NG" --transport device "$@"; }
export -f torchrun_ launch_; export NG P
[ "$NG" -ge 8 ] && export CCMPI_TEST_LARGEST_ONLY=1     # budgeted box: multi-rank tests at the largest group size only
run tests 1500 python -m pytest tests --maxfail=3 -q -m gpu
run bench_ours 400 bash -c 'torchrun_ bench.py --gpus $NG --steps 20 --warmup 5'
run bench_nccl 400 bash -c 'torchrun_ bench.py --impl nccl --gpus $NG --steps 20 --warmup 5'
run sweep 900 bash -c 'launch_ python benchmarks/sweep.py --tight --min-kb 1 --max-mb ${SWEEP_MAX_MB:-256}'
run fused 600 bash -c 'launch_ python benchmarks/fused_bench.py'
run seq 600 bash -c 'launch_ python benchmarks/seq_bench.py'
run e2e_probe 300 python scripts/e2e_probe.py
run profile 300 bash -c 'torchrun_ benchmarks/profile_step.py ours'
if [ "$NG" -ge 8 ]; then
    # the N = 4 point of the scaling curve on the first four GPUs, both GPU arms
    export NG4=4
    run bench4_ours 400 bash -c 'P=$((P + 11)); CUDA_VISIBLE_DEVICES=0,1,2,3 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $P bench.py --gpus 4 --steps 20 --warmup 5'
    run bench4_nccl 400 bash -c 'P=$((P + 12)); CUDA_VISIBLE_DEVICES=0,1,2,3 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $P bench.py --impl nccl --gpus 4 --steps 20 --warmup 5'
    # one CPU step of the reference arm at N = 8 (how long the driver's run of that arm will take)
    [ -n "$REF_PROBE" ] && run ref8_probe 400 bash -c 'P=$((P + 13)); CCMPI_REF_BUDGET_S=${REF_PROBE_BUDGET_S:-60} python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $P bench.py --impl reference --gpus 8 --steps 2 --warmup 3'
fi
[ -n "$EXTRA" ] && run extra 900 bash -c "$EXTRA"
cat "$OUT"
