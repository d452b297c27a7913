#!/bin/bash
# Budgeted multi-GPU session for the numbers of record (4 or 8 GPUs): selected multi-rank tests, the headline
# bench on all GPUs of the box, the collective sweep, and -- on 8 GPUs -- configs 1, 3 and 4.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
export CCMPI_DEVICE_TIMEOUT_S=${CCMPI_DEVICE_TIMEOUT_S:-20}
export CCMPI_TEST_LARGEST_ONLY=1
OUT=gpurun_out/final_${NG}gpu.txt
: > "$OUT"
P=29800
run() { local label=$1 limit=$2; shift 2
    case " $SKIP " in *" $label "*) echo "[$label] skipped" | tee -a "$OUT"; return;; esac
    local t0=$(date +%s)
    timeout "$limit" "$@" > gpurun_out/f${NG}_${label}.log 2>&1
    echo "[$label] rc=$? $(( $(date +%s) - t0 ))s  $(tail -1 gpurun_out/f${NG}_${label}.log | cut -c1-700)" | tee -a "$OUT"; }
torchrun_() { P=$((P + 1)); python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NG" --master-addr 127.0.0.1 --master-port $P "$@"; }
launch_() { python -m collective_communication_mpi_b200.launch -n "$NG" --transport device "$@"; }
export -f torchrun_ launch_; export NG P
run tests 600 python -m pytest tests/test_gpu_seq.py tests/test_gpu_multirank.py tests/test_gpu_p2p_device.py -q --maxfail=4 \
    -k "sequence_parallel_block or megatron_layout or dp_allreduce_fused or match_dense_model or travel_over or collectives_on_device_plane"
run bench_ours 400 bash -c 'torchrun_ bench.py --gpus $NG --steps 20 --warmup 5'
[ -n "$WITH_NCCL" ] && run bench_nccl 400 bash -c 'torchrun_ bench.py --impl nccl --gpus $NG --steps 20 --warmup 5'
run sweep 900 bash -c 'launch_ python benchmarks/sweep.py --tight --min-kb 1 --max-mb ${SWEEP_MAX_MB:-1024}'
if [ "$NG" -ge 8 ]; then
    run config1 300 python scripts/config1.py -n 8 --out gpurun_out/config1_8gpu_box.jsonl
    run mx_bench 400 bash -c 'launch_ python benchmarks/mx_bench.py'
    run config_tp8 400 bash -c 'launch_ python benchmarks/config_bench.py --config tp8'
    run config_dp8 400 bash -c 'launch_ python benchmarks/config_bench.py --config dp8'
fi
[ -n "$EXTRA" ] && run extra 900 bash -c "$EXTRA"
cat "$OUT"
