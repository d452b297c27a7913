#!/bin/bash
# What has not run on hardware yet (written or fixed after round 2's GPU minutes were spent), each
# under its own timeout. Needs >= 2 GPUs for the first item, 8 for the last.
#   gpurun --gpus 2 --timeout 900  -- 'bash scripts/gpu_next.sh'
#   gpurun --gpus 8 --timeout 1200 -- 'bash scripts/gpu_next.sh'
# Results: gpurun_out/next_<N>gpu.txt and one log per item.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
OUT=gpurun_out/next_${NG}gpu.txt
: > "$OUT"
run() {   # run <label> <timeout> <command...>
    local label=$1 limit=$2; shift 2
    local log=gpurun_out/next_${label}_${NG}gpu.log
    timeout "$limit" "$@" > "$log" 2>&1
    local rc=$?
    echo "[$label] rc=$rc  $(tail -1 "$log")" | tee -a "$OUT"
}
# 1. closing fused ops: flag words zeroed between barriers and reused by another kind of op
run slot_reuse 600 python -m pytest tests/test_zzz_slot_reuse.py -q     # close() / set_tokens() / shared scatter op of a block stack
# 1b. one flag region, launches with changing CTA counts (absolute ready-flag targets)
run flag_targets 300 python -m pytest tests/test_zzz_flag_targets.py -q
# 2. smoke() as the driver runs it (five steps, loss must fall)
run smoke 300 python -c 'import __graft_entry__ as g; g.smoke(); print("SMOKE OK")'
if [ "$NG" -ge 8 ]; then
    # 3. BASELINE config 4 (dp8: weights replicated, gradient all-reduce fused with the update) after the
    #    last-bucket ordering fix in parallel/dp.py; mp2 x dp4 was re-measured after the fix, dp8 was not
    run config_dp8 400 python -m collective_communication_mpi_b200.launch -n 8 --transport device \
        python benchmarks/config_bench.py --config dp8
fi
cat "$OUT"
