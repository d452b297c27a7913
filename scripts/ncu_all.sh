#!/bin/bash
# ncu --set full of one launch of every kernel family (1 GPU), after the same program has exited 0 without ncu.
# Brings back text summaries (small) and the report of the GEMM kernels with source.
mkdir -p gpurun_out
R='regex:gemm_kernel_2cta|allreduce|allgather|reduce_scatter|alltoall|barrier_kernel|gate_|mse_kernel|sgd_sr|quantize_mxfp8'
python scripts/ncu_targets.py > gpurun_out/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on --profile-from-start off -k "$R" -f -o gpurun_out/r2_prof_all \
    python scripts/ncu_targets.py > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_run.log
ncu -i gpurun_out/r2_prof_all.ncu-rep --page raw --csv > gpurun_out/r2_ncu_raw.csv 2> /dev/null
ncu -i gpurun_out/r2_prof_all.ncu-rep --page details > gpurun_out/r2_ncu_details.txt 2> /dev/null
ls -la gpurun_out/r2_prof_all.ncu-rep gpurun_out/r2_ncu_raw.csv gpurun_out/r2_ncu_details.txt
# the report itself is only brought back if it fits the 64 MiB return limit next to everything else
[ "$(stat -c %s gpurun_out/r2_prof_all.ncu-rep)" -gt 45000000 ] && rm -f gpurun_out/r2_prof_all.ncu-rep && echo "report too large to return: kept the exports"
true
