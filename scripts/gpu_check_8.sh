#!/bin/bash
# 8-GPU validation (budgeted: ~5 min wall): device-plane tests at 8 ranks, layers on 2x4 / 8x1 grids,
# bench (ours / nccl) at 8 and 4, configs tp8 / dp8, collective sweep to 1 GiB, fused-op bench at mp=8.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "== $NG GPUs"
CCMPI_TEST_LARGEST_ONLY=1 timeout 600 python -m pytest tests/test_gpu_multirank.py -q -k "True or layers_and_train_step_match_dense_model[2-4] or layers_and_train_step_match_dense_model[8-1] or layers_and_train_step_match_dense_model[4-2]" > gpurun_out/pytest_multirank_${NG}gpu.txt 2>&1; tail -3 gpurun_out/pytest_multirank_${NG}gpu.txt
port=29700
: > gpurun_out/bench_${NG}gpu_box.jsonl
for n in 8 4; do for impl in ours nccl; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --impl $impl --steps 20 --warmup 5 > /tmp/b.txt 2>&1
  grep "^{" /tmp/b.txt | tee -a gpurun_out/bench_${NG}gpu_box.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['impl'], d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])" || tail -5 /tmp/b.txt
done; done
: > gpurun_out/config_bench_${NG}gpu.jsonl
for cfg in tp8 dp8; do for impl in ours nccl; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $port benchmarks/config_bench.py --config $cfg --impl $impl > /tmp/c.txt 2>&1
  grep "^{" /tmp/c.txt | tee -a gpurun_out/config_bench_${NG}gpu.jsonl || tail -5 /tmp/c.txt
done; done
timeout 300 python -m collective_communication_mpi_b200.launch -n $NG --transport device python benchmarks/sweep.py --min-kb 4 --max-mb 1024 > gpurun_out/sweep_${NG}gpu.log 2>&1
tail -2 gpurun_out/sweep_${NG}gpu.log
FB_HIDDEN=8192 timeout 200 python -m collective_communication_mpi_b200.launch -n $NG --transport device python benchmarks/fused_bench.py 2>&1 | grep -v Warning | tail -9 | tee gpurun_out/fused_bench_${NG}gpu.txt
